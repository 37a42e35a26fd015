#!/usr/bin/env python
"""Mean CUDA-event time of the two per-cell kernels, per precision, on the paper-shape panel."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn  # noqa: E402

dev = torch.device("cuda:0")
T, N = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (240, 10000)
macro, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(T, N, 46, 178, seed=1234)]
res = {}
for prec in sys.argv[3:] or ["fp32", "bf16x3"]:
    torch.manual_seed(42)
    model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                             "b200_precision": prec}).to(dev)
    for name in ("dlap_cell_fwd", "dlap_cell_bwd"):
        for it in range(8):
            if it == 3:
                Fn.kernel_timer = Fn.KernelTimer(name)
            model.zero_grad(set_to_none=True)
            out = model(macro, x, R, mask)
            out["loss"].backward()
        torch.cuda.synchronize()
        res[f"{prec}/{name}"] = round(Fn.kernel_timer.mean_ms(), 4)
        Fn.kernel_timer = None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        model.zero_grad(set_to_none=True)
        model(macro, x, R, mask)["loss"].backward()
    e1.record(); torch.cuda.synchronize()
    res[f"{prec}/step_ms"] = round(e0.elapsed_time(e1) / 10, 4)
    res[f"{prec}/Gcells_per_s"] = round(T * N / (e0.elapsed_time(e1) / 10) / 1e6, 3)
print(json.dumps(res))
