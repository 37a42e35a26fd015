#!/usr/bin/env python
"""CUDA-event time of every phase of dlap_step_reduce on its own (paper shape), and of the whole kernel."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, _lib  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn  # noqa: E402

dev = torch.device("cuda:0")
T, N = 240, 10000
macro, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(T, N, 46, 178, seed=1234)]
torch.manual_seed(42)
model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                         "b200_precision": "bf16x3"}).to(dev).eval()
kp = [p.detach() for p in model._inputs(macro, x, True).params]
ws = Fn.workspace(T, N, 46, 178, 4, dev)
cp = Fn.cell_forward(ws, macro, x, mask, kp, None, None, "bf16x3", True)
red = Fn.step_reduce(ws, cp, R, mask, True, True)
torch.cuda.synchronize()
e = torch.empty(9, N, device=dev); t_i = torch.empty(N, device=dev)
w = torch.empty(T, N, device=dev)


def launch(lo, hi):
    _lib.call("dlap_step_reduce", ws.pp, cp.u.data_ptr(), cp.q.data_ptr(), R.data_ptr(), mask.data_ptr(),
              cp.index.data_ptr(), 1, 0, None, lo, hi, red.row_stats.data_ptr(), red.pr.data_ptr(), red.scalars.data_ptr(),
              w.data_ptr(), ws.col_part.data_ptr(), e.data_ptr(), t_i.data_ptr(), red.G.data_ptr(),
              ws.loss_part.data_ptr(), red.dS2.data_ptr(), ws.row_part.data_ptr(), ws.bar.data_ptr(), None, None,
              torch.cuda.current_stream().cuda_stream)


res = {}
only = os.environ.get("PHASES")
todo = [tuple(int(v) for v in only.split(","))] if only else [(0, 0), (1, 1), (2, 2), (3, 3), (4, 4), (0, 2), (0, 4)]
for lo, hi in todo:
    for _ in range(3):
        launch(lo, hi)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        launch(lo, hi)
    b.record(); torch.cuda.synchronize()
    res[f"phase {lo}..{hi}"] = round(a.elapsed_time(b) / 20 * 1e3, 1)
res["plan"] = {"col_splits": ws.plan.col_splits, "sms": ws.plan.sms}
print(json.dumps(res))
