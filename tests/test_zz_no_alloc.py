"""dlap_loss_finish_noalloc against dlap_loss_finish, through the C ABI.

Who needs the no-allocation entry point: hosts that may not create device memory while they run --
a service living inside a fixed, pre-partitioned slice of GPU memory; a latency-bound evaluation loop
where cudaMalloc/cudaFree are banned because they lock the driver and synchronise the device; glue
code that only receives raw pointers from another runtime.  For them the two scratch buffers of the
loss-finishing step sit in an array inside the library image (one copy per GPU, kept until exit).

The test feeds both entry points the same per-stock column sums and requires identical results
(same kernel, same summation order, so equality is exact -- tighter than the 1e-5 the parity tests
use), on two calls in a row with different inputs: the array is shared between calls and keeps the
previous call's contents, so the second call proves that nothing stale leaks through.  The file name
sorts last on purpose: with `pytest -x` every other test has already run when this one does.
"""
import ctypes as C

import pytest
import torch

from deeplearninginassetpricing_paperreplication_b200 import _lib

pytestmark = pytest.mark.gpu
KM = 8


def _plan(T, N, dev_index=0):
    dims = _lib.DlapDims(T, N, 46, 8, 4)
    plan = _lib.DlapPlan()
    _lib.call("dlap_make_plan", C.byref(dims), dev_index, C.byref(plan))
    return plan


def _both(plan, col, nsplit, dev):
    """-> ((e, t_i, scalars) from the existing entry point, same from the no-allocation one)."""
    N = plan.dims.N
    res = []
    for name in ("dlap_loss_finish", "dlap_loss_finish_noalloc"):
        e = torch.full((KM + 1, N), float("nan"), device=dev)
        t_i = torch.full((N,), float("nan"), device=dev)
        scalars = torch.zeros(8, device=dev)
        st = torch.cuda.current_stream(dev).cuda_stream
        if name == "dlap_loss_finish":
            loss_part = torch.empty(max(plan.loss_part_bytes // 8, 1), dtype=torch.float64, device=dev)
            counter = torch.zeros(1, dtype=torch.int32, device=dev)
            _lib.call(name, C.byref(plan), col.data_ptr(), nsplit, e.data_ptr(), t_i.data_ptr(),
                      loss_part.data_ptr(), counter.data_ptr(), scalars.data_ptr(), st)
        else:
            _lib.call(name, C.byref(plan), col.data_ptr(), nsplit, e.data_ptr(), t_i.data_ptr(),
                      scalars.data_ptr(), st)
        torch.cuda.synchronize(dev)
        res.append((e, t_i, scalars))
    return res


def test_noalloc_matches_existing_entry_point_twice():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(0)
    for call, (N, nsplit) in enumerate([(10_000, 5), (777, 3)]):      # second call: other size, other data
        plan = _plan(24, N)
        col = torch.randn(nsplit, KM + 2, N, generator=g, dtype=torch.float64)
        col[:, KM + 1] = torch.randint(0, 9, (nsplit, N), generator=g).double()     # month counts
        col = col.to(dev)
        (e0, t0, s0), (e1, t1, s1) = _both(plan, col, nsplit, dev)
        assert torch.isfinite(e1).all() and torch.isfinite(t1).all(), f"call {call}"
        assert torch.equal(e0, e1) and torch.equal(t0, t1), f"call {call}"
        assert torch.equal(s0[:2], s1[:2]) and float(s1[0]) > 0 and float(s1[1]) > 0, f"call {call}"


def test_noalloc_rejects_oversize_panel():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    N = 262_144 + 1                                   # DLAP_NOALLOC_MAX_STOCKS + 1
    plan = _plan(2, N)
    col = torch.zeros(1, KM + 2, N, dtype=torch.float64, device=dev)
    e = torch.zeros(KM + 1, N, device=dev); t_i = torch.zeros(N, device=dev); sc = torch.zeros(8, device=dev)
    rc = _lib.load().dlap_loss_finish_noalloc(C.byref(plan), col.data_ptr(), 1, e.data_ptr(), t_i.data_ptr(),
                                             sc.data_ptr(), None)
    assert rc == -8
    assert b"capacity" in _lib.load().dlap_error_string(rc)
    with pytest.raises(RuntimeError, match="capacity"):
        _lib.call("dlap_loss_finish_noalloc", C.byref(plan), col.data_ptr(), 1, e.data_ptr(), t_i.data_ptr(),
                  sc.data_ptr(), None)
