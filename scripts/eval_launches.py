#!/usr/bin/env python
"""The kernels of one evaluation (get_weights(normalized=True) + forward under no_grad; src/train.py:106-153) on the
paper-shape panel: run under `ncu --metrics gpu__time_duration.sum` for the launch list."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN

dev = torch.device("cuda:0")
panel = [t.to(dev) for t in bench.synthetic_panel(240, 10000, 46, 178, seed=1234)]
torch.manual_seed(42)
model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0}).to(dev).eval()
for it in range(5):
    if it == 4:
        torch.cuda.synchronize(); print("MEASURED ITERATION", flush=True)
        torch.cuda.nvtx.range_push("eval")
    with torch.no_grad():
        w, _ = model.get_weights(panel[0], panel[1], panel[3], normalized=True)
        o = model(*panel, phase="conditional")
    if it == 4:
        torch.cuda.nvtx.range_pop()
torch.cuda.synchronize()
