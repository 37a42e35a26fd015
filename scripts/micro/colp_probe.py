"""Times dlap_col_partial with pieces switched off (which input makes it slow?)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench
from deeplearninginassetpricing_paperreplication_b200 import _lib, functional as Fn
dev = torch.device("cuda:0")
T, N = 240, 10000
_, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(T, N, 46, 178, seed=1234)]
ws = Fn.workspace(T, N, 46, 178, 4, dev)
index = ws.panel_index(mask)
u = torch.randn(T, N, device=dev); g = torch.randn(8, T, N, device=dev)
rs = torch.rand(T, 8, device=dev); w = torch.empty(T, N, device=dev)
def run(name, gg, RR, uu, ww, idx):
    p = lambda t: None if t is None else t.data_ptr()
    for it in range(13):
        if it == 3:
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record()
        _lib.call("dlap_col_partial", ws.pp, p(gg), p(RR), mask.data_ptr(), p(uu), rs.data_ptr(), p(idx),
                  ws.col_part.data_ptr(), p(ww), None)
    e1.record(); torch.cuda.synchronize()
    print(f"{name:40s} {e0.elapsed_time(e1) / 10 * 1e3:7.1f} us")
run("full packed", g, R, u, w, index)
run("no w_out", g, R, u, None, index)
run("no g", None, R, u, w, index)
run("no g, no w_out", None, R, u, None, index)
run("no R", g, None, u, w, index)
run("dense (no index)", g, R, u, w, None)
run("dense, no g no w_out", None, R, u, None, None)
