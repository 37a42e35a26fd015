"""Which call invalidates a stream capture?  After every C-ABI call and a few torch ops, ask the runtime for the
capture status of the current stream."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, _lib
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
import bench

rt = ctypes.CDLL("libcudart.so.12")
def status():
    st = ctypes.c_int(-1)
    rc = rt.cudaStreamIsCapturing(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), ctypes.byref(st))
    return rc, st.value

orig = _lib.call
def traced(name, *a):
    try:
        orig(name, *a)
    finally:
        print("   ", name, "-> capture status", status(), flush=True)
_lib.call = traced
Fn._lib.call = traced

def trial(T, N, M, precision, eager_first=False, collect=False):
    print(f"== T={T} N={N} M={M} {precision} eager_first={eager_first} gc={collect}", flush=True)
    dev = torch.device("cuda:0")
    macro, x, R, mask = [t.to(dev) for t in bench.synthetic_panel(T, N, 46, M, seed=1)]
    torch.manual_seed(0)
    model = AssetPricingGAN({"macro_feature_dim": M, "individual_feature_dim": 46, "dropout": 0.0,
                             "b200_precision": precision}).to(dev).eval()
    if eager_first:
        model.zero_grad(set_to_none=True)
        out = model(macro, x, R, mask)
        out["loss"].backward()
        keep = [out[k].detach().clone() for k in ("loss", "weights")]
        del out
        torch.cuda.synchronize()
    if collect:
        import gc; gc.collect()
    try:
        gs = model.graphed_step(macro, x, R, mask)
        gs.step(); torch.cuda.synchronize()
        print("   OK, launches", gs.launches, "loss", float(gs.out["loss"]), flush=True)
    except Exception as e:
        print("   FAILED:", repr(e)[:300], flush=True)
        torch.cuda.synchronize()

_lib.call = orig; Fn._lib.call = orig
for args in [(12, 37, 178, "fp32", True), (12, 37, 178, "fp32", True, True), (12, 37, 178, "bf16x3", True),
             (240, 10000, 178, "bf16x3", True), (240, 10000, 178, "bf16x3", True, True)]:
    trial(*args)
