#!/usr/bin/env python
"""Per-phase cycle counts of the tensor-core backward kernel (CTA 0), from dlap_debug_timeline stamps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, _lib

dev = torch.device("cuda:0")
macro, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(240, 10000, 46, 178, seed=1234)]
torch.manual_seed(42)
model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                         "b200_precision": "bf16x3"}).to(dev)
for _ in range(3):
    model.zero_grad(); model(macro, x, R, mask)["loss"].backward()
st = torch.zeros(256 * 16, dtype=torch.int64, device=dev)
_lib.call("dlap_debug_timeline", st.data_ptr())
model.zero_grad(); model(macro, x, R, mask)["loss"].backward()
torch.cuda.synchronize()
_lib.call("dlap_debug_timeline", None)
s = st.view(256, 16).cpu().double()
lo, hi = 10, 110
def d(a, b, shift=0):
    return float((s[lo + shift:hi + shift, b] - s[lo:hi, a]).mean())
print("tile period (epilogue E2 start -> next E2 start):", d(0, 0, 1))
print("epilogue: wait D2_FULL", d(0, 1), "| E2 body", d(1, 2), "| E1(i+1) incl. waits", d(2, 3),
      "| wait D3_FULL", d(3, 4), "| E3 body", d(4, 5), "| tail (flush etc.)", d(5, 0, 1))
print("MMA: wait DZ2_FULL -> issued WG2+DG", d(8, 9), "| L2(i+1) incl. wait h1", d(9, 10), "| wait DZ1_FULL", d(10, 11),
      "| WG1 + L1(i+2) issue", d(11, 12), "| next wait DZ2_FULL", d(12, 8, 1))
