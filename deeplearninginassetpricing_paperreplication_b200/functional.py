"""The adversarial-SDF step as one ``torch.autograd.Function`` over the C ABI.

Forward  : macro_fwd -> [panel_index] -> cell_fwd -> row_finish -> col_partial -> loss_finish
Backward : loss_coeff -> ds_rows -> cell_bwd -> grad_finish -> macro_bwd

With the tensor-core precision the LSTM recurrence of macro_fwd (serial, one warp) runs on a side stream BESIDE
cell_fwd, which waits month by month for the per-month biases it needs (``overlap_macro``).

Every stage is a kernel launch through ``libdlap_b200.so`` on torch's current stream; torch is
used for device memory and streams only.  Nothing here synchronises with the host, so a whole
training step can be captured in a CUDA graph.

Month-sharded multi-GPU runs (``shard`` argument): each rank owns a slab of months, or every n-th month,
of ``x``/``R``/``mask`` and the full macro series.  The only exchanges are an all-reduce of the
per-stock column sums ``[10, N]`` in the forward pass and, in the backward pass, an all-reduce
of the per-month ``da_t | dc_t`` rows (zero outside the rank's slab) together with the
per-cell weight gradients; every rank then runs the (tiny, serial) LSTM BPTT redundantly.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import DlapDims, DlapPlan, DlapTensors

RS_W = 8      # row_stats row width  (csrc/dlap_common.cuh)
SC_W = 8      # scalars width
SC_LC, SC_LU, SC_SHARPE, SC_NBAR = 0, 1, 2, 3
RS_NT, RS_MU, RS_P, RS_PR, RS_S, RS_RBAR = 0, 1, 2, 3, 4, 5
KM = 8
TC_MAX_F = 46   # characteristics the tensor-core kernels take (the ones column of the X image sits at k = 46)

_PRECISIONS = {"fp32": _lib.DLAP_PREC_FP32, "bf16x3": _lib.DLAP_PREC_BF16X3}
_BWD_PRECISION = {"fp32": "fp32", "bf16x3": "bf16x3"}     # arithmetic of dlap_cell_bwd per model precision


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream(dev: torch.device):
    return torch.cuda.current_stream(dev).cuda_stream


def _tensors(ts) -> DlapTensors:
    s = DlapTensors()
    for name, t in zip(DlapTensors.FIELDS, ts):
        setattr(s, name, _ptr(t))
    return s


class KernelTimer:
    """CUDA-event bracket around every launch of one entry point (bench.py's roofline leg): the
    events are recorded on the launching stream, read after the caller synchronises."""

    def __init__(self, name: str):
        self.name, self.pairs = name, []

    def bracket(self, name, dev):
        if name != self.name:
            return None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(torch.cuda.current_stream(dev))
        self.pairs.append((e0, e1))
        return e1

    def mean_ms(self) -> float:
        ts = [a.elapsed_time(b) for a, b in self.pairs]
        return sum(ts) / max(len(ts), 1)


kernel_timer: Optional[KernelTimer] = None


def _timed_call(name, dev, *args):
    end = kernel_timer.bracket(name, dev) if kernel_timer is not None else None
    _lib.call(name, *args)
    if end is not None:
        end.record(torch.cuda.current_stream(dev))


class _timed_region:
    """`with _timed_region(name, dev):` -- the enclosed stream work is bracketed like an entry point."""

    def __init__(self, name, dev):
        self.name, self.dev = name, dev

    def __enter__(self):
        self.end = kernel_timer.bracket(self.name, self.dev) if kernel_timer is not None else None

    def __exit__(self, *exc):
        if self.end is not None:
            self.end.record(torch.cuda.current_stream(self.dev))
        return False


class Workspace:
    """Plan + pure scratch buffers for one (shape, device).  Scratch is reused by every call on
    that shape (single-stream semantics, like the reference); nothing that the backward pass
    needs from the forward pass lives here."""

    def __init__(self, T: int, N: int, F: int, M: int, H: int, device: torch.device, reserve: int = 0):
        self.device = device
        dims = DlapDims(T, N, F, M, H if M > 0 else 0)
        self.plan = DlapPlan()
        _lib.call("dlap_make_plan", C.byref(dims), device.index if device.index is not None
                  else torch.cuda.current_device(), C.byref(self.plan))
        p = self.plan
        # SMs left to a concurrent dlap_panel_pull (reserve_sms): the persistent per-cell kernels take one CTA
        # per SM and all of its registers, so they only overlap with the ingest if they leave some SMs alone
        self.reserve = min(int(reserve), p.grid_cell - 1) if reserve > 0 else 0
        p.grid_cell -= self.reserve
        self.pp = C.byref(p)
        f64 = lambda nbytes: torch.empty(max(nbytes // 8, 1), dtype=torch.float64, device=device)
        f32 = lambda nbytes: torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=device)
        self.row_part = f64(p.row_part_bytes)
        self.col_part = f64(p.col_part_bytes)
        self.loss_part = f64(p.loss_part_bytes)
        self.ds_part = f64(p.ds_part_bytes)
        self.wpart = f32(p.wpart_bytes)
        self.dadc_part = f32(p.dadc_part_bytes)
        self.dpre = f32(p.dpre_bytes)
        self.counter = torch.zeros(1, dtype=torch.int32, device=device)
        self.months_ready = torch.zeros(1, dtype=torch.int32, device=device)   # progress of a concurrent dlap_macro_fwd
        self.side = torch.cuda.Stream(device)                                   # the stream it runs on
        self.launches = 0          # kernels launched through this workspace (bench bookkeeping)

    def panel_index(self, mask: torch.Tensor) -> torch.Tensor:
        return panel_index(self, mask)


_index_cache = []          # up to 4 x (weakref to a mask tensor, key, index, build event, build stream), most recent last


def panel_index(ws: "Workspace", mask: torch.Tensor) -> torch.Tensor:
    """dlap_panel_index for `mask` (tensor-core kernels, dlap_panel_pull).  Reused only while the caller passes
    the very same tensor object with an unchanged version counter (full-batch training loops do; a loop that
    double-buffers its panels alternates between two); any other mask gets a freshly allocated index, so an
    index captured for a pending backward is never overwritten.  The index depends on T and N only, so one
    small cache serves every workspace.
    The key is (tensor object, version counter, data pointer, device).  Writes that bypass the version counter
    (``mask.data = ...``, DLPack / CUDA-array-interface aliases, foreign kernels) are invisible to it: call
    ``invalidate_panel_index()`` after such a write.  The index is built on the stream that is current at the
    time; a later call on ANOTHER stream waits for the build event before it hands the index out."""
    cur = torch.cuda.current_stream(ws.device)
    key = (mask._version, mask.data_ptr(), mask.device)
    for k, hit in enumerate(_index_cache):
        if hit[0]() is mask and hit[1] == key:
            _index_cache.append(_index_cache.pop(k))
            if hit[4] != cur.cuda_stream and not torch.cuda.is_current_stream_capturing():
                cur.wait_event(hit[3])
            return hit[2]
    index = torch.empty(max(ws.plan.index_bytes // 4, 4), dtype=torch.int32, device=ws.device)
    _lib.call("dlap_panel_index", ws.pp, _ptr(mask), _ptr(index), cur.cuda_stream)
    ws.launches += 2
    built = torch.cuda.Event()
    if not torch.cuda.is_current_stream_capturing():
        built.record(cur)
    _index_cache.append((weakref.ref(mask), key, index, built, cur.cuda_stream))
    del _index_cache[:-4]
    return index


def invalidate_panel_index() -> None:
    """Forget every cached valid-cell index (see panel_index)."""
    del _index_cache[:]


_workspaces: Dict[Tuple, Workspace] = {}

# SMs that the per-cell kernels leave free so that upload_panel's pull kernel of the NEXT panel can run beside
# them (a training loop that double-buffers its panels sets this before its first step; 0 = use every SM)
reserve_sms = 0


def workspace(T, N, F, M, H, device) -> Workspace:
    key = (T, N, F, M, H if M > 0 else 0, device.index, reserve_sms)
    ws = _workspaces.get(key)
    if ws is None:
        ws = _workspaces[key] = Workspace(T, N, F, M, H, device, reserve_sms)
    return ws


@dataclass
class StepOptions:
    weighted_loss: bool = True
    need_moments: bool = True       # False in phase 'unconditional': the moment net is unused
    precision: str = "fp32"
    dropout_p: float = 0.0          # > 0 only in train mode (src/model.py:213)
    dropout_seed: int = 0
    dropout_seed_dev: Optional[torch.Tensor] = None   # int64 [1] on the device: overrides dropout_seed (GraphedStep)
    shard: Optional["MonthShard"] = None


@dataclass
class MonthShard:
    """This rank's months t0, t0 + stride, ... (< t1) of a T_total-month panel (see module docstring): a slab
    (stride 1) or every stride-th month.  Dealing the months round-robin (rank r of n: t0 = r, stride = n) lets every
    rank follow the LSTM recurrence from the first month on, instead of waiting until it reaches the rank's slab."""
    group: object
    t0: int
    t1: int
    T_total: int
    n_bar: torch.Tensor             # [1] float32, global mean of clamp(N_t, 1)
    stride: int = 1

    def rows(self, table: torch.Tensor) -> torch.Tensor:
        """This rank's rows of a per-month table of the whole series (a view)."""
        return table[self.t0:self.t1:self.stride]


def make_month_shard(mask_local: torch.Tensor, t0: int, T_total: int, group=None, stride: int = 1) -> MonthShard:
    """MonthShard for this rank's months `mask_local` = months t0, t0 + stride, ... of a T_total-month panel.
    The only mask-derived global quantity the row stage needs is N_bar = mean_t clamp(N_t, 1)
    (src/model.py:359-360); it is all-reduced here, once per panel, not per step."""
    import torch.distributed as dist
    n_t = mask_local.sum(dim=1).clamp(min=1).to(torch.float32).sum().reshape(1)
    dist.all_reduce(n_t, group=group)
    n_bar = (n_t / float(T_total)).to(torch.float32).contiguous()
    t1 = int(t0) + (int(mask_local.shape[0]) - 1) * int(stride) + 1
    if stride < 1 or t0 < 0 or t1 > T_total:
        raise ValueError("month shard outside the series")
    return MonthShard(group=group, t0=int(t0), t1=t1, T_total=int(T_total), n_bar=n_bar, stride=int(stride))


def upload_panel(macro, x, R, mask, device, out=None, clear=True):
    """`batch = {k: v.to(device) ...}` of src/train.py:69-75 for one panel: (macro, x, R, mask) on `device`,
    enqueued on the current stream.  When x sits in pinned host memory only the rows of unmasked stock-months
    cross the bus (dlap_panel_pull: the kernel reads them straight out of host memory; masked rows arrive as
    zeros, which is what the loader stores there anyway); otherwise x is copied whole.  `out`: device tensors
    to fill instead of allocating (a training loop that double-buffers its panels).  clear=False: the caller has
    already zeroed out[1] (e.g. right after the step that used it, when the GPU has SMs to spare) or does not need
    zeros in the masked rows; the pull then starts without a memset of the whole panel in front of it."""
    device = torch.device(device)
    host = (macro, x, R, mask)
    if out is None:
        out = [None if t is None else torch.empty_like(t, device=device) for t in host]
    d_macro, d_x, d_R, d_mask = out
    with torch.cuda.device(device):
        d_mask.copy_(mask, non_blocking=True)
        if R is not None:
            d_R.copy_(R, non_blocking=True)
        if macro is not None:
            d_macro.copy_(macro, non_blocking=True)
        T, N, F = x.shape
        if x.is_pinned() and x.dtype == torch.float32 and x.is_contiguous() and mask.dtype == torch.bool:
            ws = workspace(T, N, F, 0, 0, device)
            index = panel_index(ws, d_mask)
            _lib.call("dlap_panel_pull", ws.pp, x.data_ptr(), _ptr(index), _ptr(d_x), int(reserve_sms),
                      1 if clear else 0, _stream(device))
            ws.launches += 1
        else:
            d_x.copy_(x, non_blocking=True)
    return d_macro, d_x, d_R, d_mask


def check_panel(macro, x, R, mask) -> None:
    """The reference needs contiguous inputs too (x.view at src/model.py:143,261); beyond that
    the kernels need fp32 CUDA tensors.  No silent conversion, no CPU fallback."""
    if not x.is_cuda:
        raise ValueError("individual_features must be a CUDA tensor: this build has no CPU path")
    if x.dtype != torch.float32 or x.dim() != 3 or not x.is_contiguous():
        raise ValueError("individual_features must be a contiguous float32 [T,N,F] tensor")
    T, N, _ = x.shape
    if mask.dtype != torch.bool or tuple(mask.shape) != (T, N) or not mask.is_contiguous() \
            or mask.device != x.device:
        raise ValueError("mask must be a contiguous bool [T,N] tensor on the same device")
    if R is not None and (R.dtype != torch.float32 or tuple(R.shape) != (T, N)
                          or not R.is_contiguous() or R.device != x.device):
        raise ValueError("returns must be a contiguous float32 [T,N] tensor on the same device")
    if macro is not None and (macro.dtype != torch.float32 or macro.dim() != 2
                              or not macro.is_contiguous() or macro.device != x.device):
        raise ValueError("macro_features must be a contiguous float32 [T,M] tensor on the same device")


def _flat_state(h, H, dev):
    if h is None:
        return None
    h = h.detach().reshape(-1).to(device=dev, dtype=torch.float32).contiguous()
    if h.numel() != H:
        raise ValueError(f"hidden state must have {H} elements")
    return h


# bf16x3 only: dlap_macro_fwd on a side stream, dlap_cell_fwd beside it (False, or DLAP_OVERLAP_MACRO=0 in the
# environment: one after the other)
overlap_macro = os.environ.get("DLAP_OVERLAP_MACRO", "1") != "0"


def macro_forward(ws: Workspace, macro, params, h0, c0, concurrent: bool = False):
    """dlap_macro_fwd: LSTM states + per-month biases.  `macro` covers all months of the panel.
    concurrent: enqueue it on ws.side behind the work already on the current stream and publish its progress in
    ws.months_ready; the caller hands that counter to the dlap_cell_fwd it enqueues next on the current stream and
    then joins ws.side (`join_macro`)."""
    dev = ws.device
    T, H = ws.plan.dims.T, max(ws.plan.dims.H, 1)
    save = torch.empty(T, 6 * H, dtype=torch.float32, device=dev)
    a_t = torch.empty(T, 64, dtype=torch.float32, device=dev)
    c_t = torch.empty(T, KM, dtype=torch.float32, device=dev)
    hn = torch.empty(H, dtype=torch.float32, device=dev)
    cn = torch.empty(H, dtype=torch.float32, device=dev)
    tens = _tensors(params)
    if concurrent:
        main = torch.cuda.current_stream(dev)
        ws.months_ready.zero_()                    # ordered before both the producer and the consumer
        ws.side.wait_stream(main)
        _lib.call("dlap_macro_fwd", ws.pp, _ptr(macro), C.byref(tens), _ptr(h0), _ptr(c0), _ptr(save),
                  _ptr(a_t), _ptr(c_t), _ptr(hn), _ptr(cn), _ptr(ws.dpre), _ptr(ws.months_ready),
                  ws.side.cuda_stream)
    else:
        _timed_call("dlap_macro_fwd", dev, ws.pp, _ptr(macro), C.byref(tens), _ptr(h0), _ptr(c0), _ptr(save),
                    _ptr(a_t), _ptr(c_t), _ptr(hn), _ptr(cn), _ptr(ws.dpre), None, _stream(dev))
    ws.launches += 1
    return save, a_t, c_t, hn, cn


def join_macro(ws: Workspace) -> None:
    """The current stream continues only after the concurrent dlap_macro_fwd on ws.side has finished."""
    torch.cuda.current_stream(ws.device).wait_stream(ws.side)


class GanStepFn(torch.autograd.Function):
    """(L_c, L_u, pr, weights, sharpe, h_n, c_n) = step(panel, params)."""

    @staticmethod
    def forward(ctx, opts: StepOptions, macro, x, R, mask, h0, c0, *params):
        dev = x.device
        T, N, F = x.shape
        M = 0 if macro is None else macro.shape[1]
        H = 0 if macro is None else params[1].shape[1]
        prec = _PRECISIONS[opts.precision]
        sh = opts.shard
        with torch.cuda.device(dev):
            st = _stream(dev)
            ws = workspace(T, N, F, M, H, dev)
            tens = _tensors(params)
            index = ws.panel_index(mask) if opts.precision == "bf16x3" else None
            # not while SMs are set aside for a concurrent panel pull: measured slower there (the recurrence's CTA
            # ends up beside a pull CTA, and the step is bound by the bus anyway)
            beside = overlap_macro and opts.precision == "bf16x3" and M > 0 and ws.reserve == 0
            if sh is None:
                wsm, t_base, t_stride = ws, 0, 1
                save, a_t, c_t, hn, cn = macro_forward(ws, macro, params, h0, c0, beside)
                a_loc, c_loc = a_t, c_t
            else:   # LSTM replicated over the full series; this rank uses its rows of a_t / c_t
                wsm, t_base, t_stride = workspace(sh.T_total, N, F, M, H, dev), sh.t0, sh.stride
                save, a_t, c_t, hn, cn = macro_forward(wsm, macro, params, h0, c0, beside)
                if opts.precision == "bf16x3":      # the kernel steps through the full tables itself
                    a_loc, c_loc = a_t[sh.t0:], c_t[sh.t0:]
                else:
                    a_loc, c_loc, t_stride = sh.rows(a_t).contiguous(), sh.rows(c_t).contiguous(), 1
            u = torch.empty(T, N, dtype=torch.float32, device=dev)
            g = torch.empty(KM, T, N, dtype=torch.float32, device=dev) if opts.need_moments else None
            _timed_call("dlap_cell_fwd", dev, ws.pp, _ptr(x), _ptr(R), _ptr(mask), _ptr(index), C.byref(tens), _ptr(a_loc),
                      _ptr(c_loc), _ptr(u), _ptr(g), _ptr(ws.row_part), prec, float(opts.dropout_p),
                      int(opts.dropout_seed), _ptr(opts.dropout_seed_dev),
                      _ptr(wsm.months_ready) if beside else None, t_base, t_stride, st)
            if beside:
                join_macro(wsm)
            row_stats = torch.empty(T, RS_W, dtype=torch.float32, device=dev)
            pr = torch.empty(T, dtype=torch.float32, device=dev)
            scalars = torch.empty(SC_W, dtype=torch.float32, device=dev)
            _timed_call("dlap_row_finish", dev, ws.pp, _ptr(ws.row_part), _ptr(index), prec,
                      int(opts.weighted_loss),
                      _ptr(sh.n_bar) if sh is not None else None, _ptr(row_stats), _ptr(pr),
                      _ptr(scalars), st)
            w = torch.empty(T, N, dtype=torch.float32, device=dev)
            _timed_call("dlap_col_partial", dev, ws.pp, _ptr(g), _ptr(R), _ptr(mask), _ptr(u), _ptr(row_stats),
                      _ptr(index), _ptr(ws.col_part), _ptr(w), st)
            e = torch.empty(KM + 1, N, dtype=torch.float32, device=dev)
            t_i = torch.empty(N, dtype=torch.float32, device=dev)
            if sh is None:
                col, nsplit = ws.col_part, ws.plan.col_splits
            else:   # exchange 1: per-stock sums over ALL months
                with _timed_region("exchange_fwd", dev):
                    col = ws.col_part.view(ws.plan.col_splits, KM + 2, N)
                    col = col[0] if ws.plan.col_splits == 1 else col.sum(dim=0)   # one split: reduce it in place
                    torch.distributed.all_reduce(col, group=sh.group)
                nsplit = 1
            _timed_call("dlap_loss_finish", dev, ws.pp, _ptr(col), nsplit, _ptr(e), _ptr(t_i),
                      _ptr(ws.loss_part), _ptr(ws.counter), _ptr(scalars), st)
            ws.launches += 5
        ctx.set_materialize_grads(False)      # unused outputs come back as None, not as zero tensors (fill kernels)
        ctx.save_for_backward(x, R, mask, macro, h0, c0, *params)
        ctx.opts, ctx.ws = opts, ws
        ctx.inter = (save, a_t, c_t, row_stats, scalars, g, e, t_i, index)
        l_c, l_u, sharpe = scalars[SC_LC], scalars[SC_LU], scalars[SC_SHARPE]
        ctx.mark_non_differentiable(w, sharpe, hn, cn)
        return l_c, l_u, pr, w, sharpe, hn, cn

    @staticmethod
    def backward(ctx, g_lc, g_lu, g_pr, *_unused):
        x, R, mask, macro, h0, c0, *params = ctx.saved_tensors
        save, a_t, c_t, row_stats, scalars, g, e, t_i, index = ctx.inter
        opts, ws = ctx.opts, ctx.ws
        sh = opts.shard
        dev = x.device
        T, N, F = x.shape
        M = 0 if macro is None else macro.shape[1]
        need = ctx.needs_input_grad[7:]
        want_sdf = any(need[:10])
        want_mom = any(need[10:]) and g is not None
        if not (want_sdf or want_mom):
            return (None,) * (7 + len(params))
        prec = _PRECISIONS[_BWD_PRECISION[opts.precision]]

        def scalar(t):
            return None if t is None else t.detach().reshape(1).to(torch.float32).contiguous()

        g_lc, g_lu = scalar(g_lc), scalar(g_lu)
        if g_pr is not None:
            g_pr = g_pr.detach().to(torch.float32).contiguous()
        with torch.cuda.device(dev):
            st = _stream(dev)
            tens = _tensors(params)
            G = torch.empty(KM + 1, N, dtype=torch.float32, device=dev)
            _timed_call("dlap_loss_coeff", dev, ws.pp, _ptr(e), _ptr(t_i), _ptr(g_lc) if g is not None else None,
                      _ptr(g_lu), _ptr(G), st)
            _timed_call("dlap_ds_rows", dev, ws.pp, _ptr(g), _ptr(R), _ptr(mask), _ptr(G), _ptr(index), _ptr(ws.ds_part), st)
            if sh is None:
                a_loc, c_loc = a_t, c_t
            else:
                a_loc, c_loc = sh.rows(a_t).contiguous(), sh.rows(c_t).contiguous()
            _timed_call("dlap_cell_bwd", dev, ws.pp, _ptr(x), _ptr(R), _ptr(mask), _ptr(index), C.byref(tens), _ptr(a_loc),
                      _ptr(c_loc), _ptr(row_stats), _ptr(ws.ds_part), _ptr(g_pr), _ptr(G), _ptr(scalars),
                      int(opts.weighted_loss), int(want_sdf), int(want_mom), _ptr(ws.wpart),
                      _ptr(ws.dadc_part), prec, float(opts.dropout_p), int(opts.dropout_seed),
                      _ptr(opts.dropout_seed_dev), st)
            # All twelve gradient tensors are views of one flat buffer.  Month-sharded: the buffer starts with the
            # [T_total,72] da_t | dc_t rows (this rank fills its slab) and is zero-filled, so that exchange 2 is ONE
            # all-reduce over it -- gradients that are computed later (LSTM, b1, bq) are still zero.  Otherwise no
            # fill: dlap_grad_finish and dlap_macro_bwd write every element of every gradient that is handed back
            # (a fill is one more kernel in a chain of short dependent ones).
            sizes = [0 if p is None else p.numel() for p in params]
            head = 0 if sh is None else sh.T_total * 72
            buf = (torch.empty if sh is None else torch.zeros)(head + sum(sizes), dtype=torch.float32, device=dev)
            flat_g = buf[head:]
            grads, off = [], 0
            for p, n in zip(params, sizes):
                grads.append(None if p is None else flat_g[off:off + n].view_as(p))
                off += n
            gt = _tensors(grads)
            if sh is None:
                dadc_all = dadc = torch.empty(T, 72, dtype=torch.float32, device=dev)
            else:
                dadc_all = buf[:head].view(sh.T_total, 72)
                dadc = sh.rows(dadc_all) if sh.stride == 1 else torch.empty(T, 72, dtype=torch.float32, device=dev)
            _timed_call("dlap_grad_finish", dev, ws.pp, _ptr(ws.wpart), _ptr(ws.dadc_part), _ptr(index), prec, int(want_sdf),
                      int(want_mom), C.byref(gt), _ptr(dadc), st)
            ws.launches += 4
            if sh is not None and sh.stride != 1:
                sh.rows(dadc_all).copy_(dadc)
            if sh is None:
                wsm = ws
            else:   # exchange 2: per-cell weight grads + every month's da_t | dc_t row
                wsm = workspace(sh.T_total, N, F, M, ws.plan.dims.H, dev)
                with _timed_region("exchange_bwd", dev):
                    torch.distributed.all_reduce(buf, group=sh.group)
            _timed_call("dlap_macro_bwd", dev, wsm.pp, _ptr(macro), C.byref(tens), _ptr(h0), _ptr(c0), _ptr(save),
                      _ptr(dadc_all), _ptr(wsm.dpre), int(want_sdf), int(want_mom), C.byref(gt), st)
            ws.launches += 2 if (want_sdf and M > 0) else 1
        out = []
        for i, gr in enumerate(grads):
            live = need[i] and ((i < 10 and want_sdf) or (i >= 10 and want_mom))
            out.append(gr if live else None)
        return (None,) * 7 + tuple(out)


def sdf_weights(macro, x, mask, params, h0=None, c0=None, normalized=False, precision="fp32",
                weighted_loss=True):
    """Inference path of SDFNetwork.forward / get_weights (src/model.py:221-281, :565-594):
    weights [T,N] (demeaned, optionally |w|-normalised per month) and the final LSTM state."""
    dev = x.device
    T, N, F = x.shape
    M = 0 if macro is None else macro.shape[1]
    H = 0 if macro is None else params[1].shape[1]
    with torch.cuda.device(dev):
        st = _stream(dev)
        ws = workspace(T, N, F, M, H, dev)
        tens = _tensors(params)
        index = ws.panel_index(mask) if precision == "bf16x3" else None
        beside = overlap_macro and precision == "bf16x3" and M > 0 and ws.reserve == 0
        save, a_t, c_t, hn, cn = macro_forward(ws, macro, params, h0, c0, beside)
        u = torch.empty(T, N, dtype=torch.float32, device=dev)
        _lib.call("dlap_cell_fwd", ws.pp, _ptr(x), None, _ptr(mask), _ptr(index), C.byref(tens), _ptr(a_t), _ptr(c_t),
                  _ptr(u), None, _ptr(ws.row_part), _PRECISIONS[precision], 0.0, 0, None,
                  _ptr(ws.months_ready) if beside else None, 0, 1, st)
        if beside:
            join_macro(ws)
        row_stats = torch.empty(T, RS_W, dtype=torch.float32, device=dev)
        pr = torch.empty(T, dtype=torch.float32, device=dev)
        scalars = torch.empty(SC_W, dtype=torch.float32, device=dev)
        _timed_call("dlap_row_finish", dev, ws.pp, _ptr(ws.row_part), _ptr(index), _PRECISIONS[precision], int(weighted_loss), None,
                  _ptr(row_stats),
                  _ptr(pr), _ptr(scalars), st)
        w = torch.empty(T, N, dtype=torch.float32, device=dev)
        _lib.call("dlap_weights", ws.pp, _ptr(u), _ptr(mask), _ptr(row_stats), _ptr(index), int(normalized), _ptr(w), st)
        ws.launches += 4
    return w, hn, cn, save
