#!/usr/bin/env python
"""Warm CUDA-event time of every C-ABI entry point inside the fwd+bwd step (paper shape, bf16x3 by default)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn  # noqa: E402

dev = torch.device("cuda:0")
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
macro, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(240, 10000, 46, 178, seed=1234)]
torch.manual_seed(42)
model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                         "b200_precision": prec}).to(dev)
names = ["dlap_macro_fwd", "dlap_cell_fwd", "dlap_row_finish", "dlap_col_partial", "dlap_loss_finish",
         "dlap_loss_coeff", "dlap_ds_rows", "dlap_cell_bwd", "dlap_grad_finish", "dlap_macro_bwd"]
res = {}
for name in names:
    for it in range(23):
        if it == 3:
            Fn.kernel_timer = Fn.KernelTimer(name)
        model.zero_grad(set_to_none=True)
        model(macro, x, R, mask)["loss"].backward()
    torch.cuda.synchronize()
    res[name] = round(Fn.kernel_timer.mean_ms() * 1e3, 1)
    Fn.kernel_timer = None
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    model.zero_grad(set_to_none=True)
    model(macro, x, R, mask)["loss"].backward()
e1.record(); torch.cuda.synchronize()
res["sum_us"] = round(sum(res.values()), 1)
res["step_us"] = round(e0.elapsed_time(e1) / 50 * 1e3, 1)
print(json.dumps(res))
