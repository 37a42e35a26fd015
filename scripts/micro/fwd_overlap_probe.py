#!/usr/bin/env python
"""Forward pieces on the paper shape: dlap_macro_fwd alone, dlap_cell_fwd alone (contiguous tile order / the
block-cyclic order with the month counter already complete), and the two side by side (functional.overlap_macro)."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

import bench  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, _lib  # noqa: E402
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn  # noqa: E402

dev = torch.device("cuda", 0)
T, N = int(os.environ.get("T", 240)), int(os.environ.get("N", 10000))
macro, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(T, N, 46, 178, seed=1234)]
torch.manual_seed(42)
model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                         "b200_precision": "bf16x3"}).to(dev)
kp = [p.detach() for p in model._kernel_params(True)]
ws = Fn.workspace(T, N, 46, 178, 4, dev)
index = ws.panel_index(mask)
tens = Fn._tensors(kp)
u = torch.empty(T, N, device=dev)
g = torch.empty(8, T, N, device=dev)
st = lambda: torch.cuda.current_stream(dev).cuda_stream
state = {}


def macro_only():
    state["m"] = Fn.macro_forward(ws, macro, kp, None, None, False)


def cell(ready):
    save, a_t, c_t, hn, cn = state["m"]
    _lib.call("dlap_cell_fwd", ws.pp, x.data_ptr(), R.data_ptr(), mask.data_ptr(), index.data_ptr(), C.byref(tens),
              a_t.data_ptr(), c_t.data_ptr(), u.data_ptr(), g.data_ptr(), ws.row_part.data_ptr(), 1, 0.0, 0, None,
              ready, 0, 1, st())


def both(beside):
    state["m"] = Fn.macro_forward(ws, macro, kp, None, None, beside)
    cell(ws.months_ready.data_ptr() if beside else None)
    if beside:
        Fn.join_macro(ws)


def timed(fn, n=60):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / n * 1e3, 1)


res = {"macro_fwd": timed(macro_only)}
res["cell_fwd_contiguous"] = timed(lambda: cell(None))
ws.months_ready.fill_(T)
res["cell_fwd_cyclic_no_wait"] = timed(lambda: cell(ws.months_ready.data_ptr()))
res["one_after_the_other"] = timed(lambda: both(False))
res["side_by_side"] = timed(lambda: both(True))
print(json.dumps(res))
