"""The adversarial-SDF step as one ``torch.autograd.Function`` over the C ABI.

Forward  : [panel_index] -> cell_fwd (the macro LSTM runs on one CTA of its grid) -> step_reduce
Backward : cell_bwd -> macro_bwd (sums the per-cell partials, BPTT, month-level gradients)

Five launches per step.  Every stage is a kernel launch through ``libdlap_b200.so`` on torch's current stream;
torch is used for device memory and streams only.  Nothing here synchronises with the host, so a whole
training step can be captured in a CUDA graph.

Month-sharded multi-GPU runs (``shard`` argument): each rank owns a slab of months, or every n-th month,
of ``x``/``R``/``mask`` and the full macro series.  The only exchanges are an all-reduce of the
per-stock column sums ``[10, N]`` in the forward pass (between phases 2 and 3 of step_reduce) and, in the
backward pass, an all-reduce of the per-month ``da_t | dc_t`` rows (zero outside the rank's months) together
with the per-cell weight gradients; every rank then runs the (tiny, serial) LSTM BPTT redundantly.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from ._lib import DlapDims, DlapMacroIO, DlapPlan, DlapTensors

RS_W = 8      # row_stats row width  (csrc/dlap_common.cuh)
SC_W = 8      # scalars width
SC_LC, SC_LU, SC_SHARPE, SC_NBAR = 0, 1, 2, 3
RS_NT, RS_MU, RS_P, RS_PR, RS_S, RS_RBAR = 0, 1, 2, 3, 4, 5
KM = 8
GS = 12        # row pitch of the per-stock loss coefficients G [N, GS]
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
        # per SM and all of its registers, so they only overlap with the ingest if they leave some SMs alone; the
        # grid-synchronising kernels are then launched plainly (a cooperative grid does not share the device)
        self.reserve = min(int(reserve), p.sms - 2) if reserve > 0 else 0
        if self.reserve > 0:
            p.sms -= self.reserve
            p.grid_cell = min(p.grid_cell, p.sms)
            p.cooperative = 0
        if os.environ.get("DLAP_COOPERATIVE", "1") == "0":
            p.cooperative = 0
        self.pp = C.byref(p)
        f64 = lambda nbytes: torch.empty(max(nbytes // 8, 1), dtype=torch.float64, device=device)
        f32 = lambda nbytes: torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=device)
        self.col_part = f64(p.col_part_bytes)
        self.loss_part = f64(p.loss_part_bytes)
        self.wpart = f32(p.wpart_bytes)
        self.dadc_part = f32(p.dadc_part_bytes)
        self.dpre = f32(p.dpre_bytes)
        self.row_part = f64(p.row_part_bytes)
        self.bar = torch.zeros(max(p.bar_bytes // 4, 1), dtype=torch.int32, device=device)   # dlap_step_reduce's barrier words
        self.sync = torch.zeros(2, dtype=torch.int32, device=device)     # progress words of the fused dlap_cell_fwd
        self.counter = torch.zeros(4, dtype=torch.int32, device=device)  # dlap_macro_bwd's progress words
        self.launches = 0          # kernels launched through this workspace (bench bookkeeping)

    def panel_index(self, mask: torch.Tensor) -> torch.Tensor:
        return panel_index(self, mask)


class PlanOnly:
    """A plan without the step's scratch buffers: what the loader-side calls (dlap_panel_ingest, dlap_panel_index)
    need.  Quacks like a Workspace for `panel_index`."""

    def __init__(self, T: int, N: int, F: int, device: torch.device):
        self.device = device
        self.plan = DlapPlan()
        _lib.call("dlap_make_plan", C.byref(DlapDims(T, N, F, 0, 0)), device.index if device.index is not None
                  else torch.cuda.current_device(), C.byref(self.plan))
        self.pp = C.byref(self.plan)
        self.launches = 0


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
    """Forget every cached valid-cell index (see panel_index) and every X image built on one."""
    del _index_cache[:]
    del _image_cache[:]


# ---- X images (dlap_panel_image): the characteristics in the tensor-core kernels' operand format, once per panel ----
# DLAP_X_IMAGE=0 in the environment: never build them (the kernels gather and convert the rows of x every step; same
# results bit for bit)
use_x_image = os.environ.get("DLAP_X_IMAGE", "1") != "0"
force_x_image = False      # set by GraphedStep around its warm-up: the panel it captures is stepped on many times
IMAGE_TILE_BYTES = 24768
_image_cache = []          # up to 3 x [weakref to x, key, index, image or None, build event, build stream]


def forget_panel_image(x: torch.Tensor) -> None:
    """Drop the X image of `x` (its values were changed by something the version counter does not see)."""
    _image_cache[:] = [e for e in _image_cache if e[0]() is not None and e[0]() is not x]


def panel_image(ws: "Workspace", x: torch.Tensor, index: torch.Tensor, force: bool = False) -> Optional[torch.Tensor]:
    """The X images of panel `x` for the valid-cell `index`, or None (the kernels then gather the rows of x).
    A panel that is seen ONCE (a loop that uploads a fresh panel every step) is not worth a pass of its own: the
    image is built when the same tensor object comes back with an unchanged version counter -- the full-batch loops
    of the reference step on the same resident panel thousands of times -- or at once with `force`.  Never built
    during a stream capture.  Keyed like the valid-cell index: (tensor object, version counter, data pointer, index);
    a write that bypasses the version counter needs forget_panel_image(x)."""
    if not use_x_image or ws.plan.image_bytes == 0:
        return None
    _image_cache[:] = [e for e in _image_cache if e[0]() is not None]
    cur = torch.cuda.current_stream(ws.device)
    capturing = torch.cuda.is_current_stream_capturing()
    key = (x._version, x.data_ptr(), x.device)
    entry = None
    for k, e in enumerate(_image_cache):
        if e[0]() is x:
            if e[1] == key and e[2] is index:
                entry = e
                _image_cache.append(_image_cache.pop(k))
            else:
                del _image_cache[k]            # the panel's values or its mask changed
            break
    if entry is None:
        entry = [weakref.ref(x), key, index, None, None, None]
        _image_cache.append(entry)
        del _image_cache[:-3]
        if not force:
            return None
    if entry[3] is None:
        if capturing:
            return None
        ntiles = int(index[0].item())          # one read-back per panel: the image is sized by the tiles that exist
        image = torch.empty(max(ntiles, 1) * IMAGE_TILE_BYTES, dtype=torch.uint8, device=ws.device)
        _lib.call("dlap_panel_image", ws.pp, _ptr(x), _ptr(index), _ptr(image), cur.cuda_stream)
        ws.launches += 1
        built = torch.cuda.Event()
        built.record(cur)
        entry[3:6] = [image, built, cur.cuda_stream]
    elif entry[5] != cur.cuda_stream and not capturing:
        cur.wait_event(entry[4])
    return entry[3]


def cached_panel_image(x: torch.Tensor):
    """(index, image) of the X images currently held for `x`, or None."""
    for e in _image_cache:
        if e[0]() is x and e[3] is not None:
            return e[2], e[3]
    return None


def rebuild_panel_image(x: torch.Tensor, index: torch.Tensor, image: torch.Tensor) -> None:
    """dlap_panel_image into an existing buffer (a captured graph holds its address) after x was given new values."""
    T, N, F = x.shape
    _lib.call("dlap_panel_image", PlanOnly(T, N, F, x.device).pp, _ptr(x), _ptr(index), _ptr(image), _stream(x.device))
    for e in _image_cache:
        if e[0]() is x and e[3] is image:
            e[1] = (x._version, x.data_ptr(), x.device)


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
    shard: Optional["MonthShard"] = None            # a MonthShard or a StockShard
    demean: bool = True             # normalize_w (src/model.py:274): False = the weights are m u, not demeaned
    residual: bool = False          # residual_loss_factor > 0 (src/model.py:543-548): also return the residual loss


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


@dataclass
class StockShard:
    """This rank's slice of the STOCKS (all months) of a panel with n_total stocks.  Per-stock sums are then local
    and complete; per-month sums are partial: the forward pass all-reduces the months' four raw sums ([T,4] doubles)
    and, at its end, the two dS sums per month plus the two sums of e^2 ([T,2] + [2] doubles); the backward pass
    all-reduces the da_t | dc_t rows and the per-cell weight gradients, as month shards do.  A few kilobytes each,
    whatever the number of ranks -- and the per-stock stage does not grow with it."""
    group: object
    n_total: int
    peers: Optional["PeerExchange"] = None     # the two forward exchanges over NVLink peer memory (make_stock_shard)


class PeerExchange:
    """Exchange buffers of a stock-sharded step in memory that every rank's GPU can address (torch symmetric memory
    over NVLink / NVSwitch), for panels of up to T_max months: dlap_step_reduce then does both of its exchanges inside
    its one launch (include/dlap.h, dlap_peers) instead of three launches around two NCCL all-reduces."""

    def __init__(self, group, T_max: int, device: torch.device):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.T_max = int(T_max)
        nbytes = (6 * self.T_max + 2) * 8 + 16 * 4
        self.buf = symm.empty((nbytes + 7) // 8, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.epoch = torch.zeros(1, dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        self.handle.barrier()
        torch.cuda.synchronize(device)
        self.c = _lib.DlapPeers()
        self.c.world, self.c.rank, self.c.capacity = self.handle.world_size, self.handle.rank, self.T_max
        for r, ptr in enumerate(self.handle.buffer_ptrs):
            self.c.buf[r] = ptr
        self.c.epoch = self.epoch.data_ptr()
        # the backward pass' exchange (dlap_macro_bwd): per-cell weight gradients + every month's da_t | dc_t row
        self.gcap = 64 * 64 + 72 * 64 + 128 + 72 * self.T_max
        self.gbuf = symm.empty(2 * self.gcap + 8, dtype=torch.float32, device=device)
        self.gbuf.zero_()
        self.ghandle = symm.rendezvous(self.gbuf, group if group is not None else dist.group.WORLD)
        self.gepoch = torch.zeros(2, dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        self.ghandle.barrier()
        torch.cuda.synchronize(device)
        self.g = _lib.DlapPeers()
        self.g.world, self.g.rank, self.g.capacity = self.ghandle.world_size, self.ghandle.rank, self.gcap
        for r, ptr in enumerate(self.ghandle.buffer_ptrs):
            self.g.buf[r] = ptr
        self.g.epoch = self.gepoch.data_ptr()


def make_stock_shard(group, n_total: int, T: int, device, peer_memory: bool = True) -> StockShard:
    """StockShard for a rank that owns some of the n_total stocks of a T-month panel.  peer_memory: set up the
    in-kernel exchange over NVLink (PeerExchange) when the process group spans 2..8 GPUs that torch's symmetric
    memory can connect; otherwise, or when that fails, the exchanges are NCCL all-reduces between three launches."""
    import torch.distributed as dist
    peers = None
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if peer_memory and 2 <= world <= 8 and dist.get_backend(group) == "nccl":
        try:
            peers = PeerExchange(group, T, torch.device(device))
        except Exception as exc:        # no symmetric memory on this system: NCCL carries the exchanges
            import warnings
            warnings.warn(f"stock shards: peer-memory exchange unavailable ({exc!r}); using NCCL all-reduces")
    return StockShard(group=group, n_total=int(n_total), peers=peers)


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
            forget_panel_image(d_x)            # the kernel wrote d_x behind the version counter's back
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




# bf16x3, H = 4: the macro LSTM runs on one CTA of dlap_cell_fwd's own grid, beside the per-cell work (False, or
# DLAP_FUSE_MACRO=0 in the environment: a launch of its own in front of it; same results bit for bit)
fuse_macro = os.environ.get("DLAP_FUSE_MACRO", "1") != "0"


def macro_forward(ws: Workspace, macro, params, h0, c0):
    """dlap_macro_fwd on its own (MacroLSTM.forward as a stand-alone call): LSTM states + per-month biases of the
    plan's T months."""
    dev = ws.device
    T, H = ws.plan.dims.T, max(ws.plan.dims.H, 1)
    save = torch.empty(T, 6 * H, dtype=torch.float32, device=dev)
    a_t = torch.empty(T, 64, dtype=torch.float32, device=dev)
    c_t = torch.empty(T, KM, dtype=torch.float32, device=dev)
    hn = torch.empty(H, dtype=torch.float32, device=dev)
    cn = torch.empty(H, dtype=torch.float32, device=dev)
    tens = _tensors(params)
    _timed_call("dlap_macro_fwd", dev, ws.pp, _ptr(macro), C.byref(tens), _ptr(h0), _ptr(c0), _ptr(save),
                _ptr(a_t), _ptr(c_t), _ptr(hn), _ptr(cn), _stream(dev))
    ws.launches += 1
    return save, a_t, c_t, hn, cn


@dataclass
class CellPass:
    """What one per-cell forward pass leaves behind (dlap_cell_fwd): u, the moment outputs q (or None), the macro
    tables over the series, the final LSTM state, the valid-cell index the tensor-core kernels ran with (or None),
    and the dlap_macro_io that describes the tables to later calls."""
    u: torch.Tensor
    q: Optional[torch.Tensor]
    save: torch.Tensor
    a_t: torch.Tensor
    c_t: torch.Tensor
    hn: torch.Tensor
    cn: torch.Tensor
    index: Optional[torch.Tensor]
    mio: DlapMacroIO
    precision: str
    ximg: Optional[torch.Tensor] = None        # X images the kernels fetched their operands from (dlap_panel_image)


def cell_forward(ws: Workspace, macro, x, mask, params, h0, c0, precision: str, need_moments: bool,
                 dropout_p: float = 0.0, dropout_seed: int = 0, dropout_seed_dev=None,
                 shard: Optional["MonthShard"] = None, tables=None) -> CellPass:
    """dlap_cell_fwd: the macro LSTM over the (whole) series and both nets on every stock-month of the panel.
    tables = (a_t [T,64], c_t [T,8] or None): the per-month biases are the caller's (no LSTM in this call; the
    workspace is one without macro input)."""
    dev = x.device
    T, N, F = x.shape
    M, H = ws.plan.dims.M, max(ws.plan.dims.H, 1)
    months = shard if isinstance(shard, MonthShard) else None
    series_T = T if months is None else months.T_total
    month0, stride = (0, 1) if months is None else (months.t0, months.stride)
    index = ws.panel_index(mask) if precision == "bf16x3" else None
    ximg = panel_image(ws, x, index, force_x_image) if index is not None else None
    save = torch.empty(series_T, 6 * H, dtype=torch.float32, device=dev)
    hn = torch.empty(H, dtype=torch.float32, device=dev)
    cn = torch.empty(H, dtype=torch.float32, device=dev)
    if tables is None:
        a_t = torch.empty(series_T, 64, dtype=torch.float32, device=dev)
        c_t = torch.empty(series_T, KM, dtype=torch.float32, device=dev)
    else:
        a_t = tables[0].detach().contiguous()
        c_t = (tables[1].detach().contiguous() if tables[1] is not None
               else torch.zeros(series_T, KM, dtype=torch.float32, device=dev))
        if tuple(a_t.shape) != (series_T, 64) or tuple(c_t.shape) != (series_T, KM) or M != 0:
            raise ValueError("per-month bias tables must be [T,64] and [T,8], on a workspace without macro input")
    fused = (tables is None and fuse_macro and precision == "bf16x3" and M > 0 and ws.plan.dims.H == 4
             and ws.plan.sms >= 2)
    mio = DlapMacroIO(_ptr(macro) if (M > 0 and tables is None) else None, series_T, month0, stride,
                      0 if tables is None else 1, _ptr(h0), _ptr(c0), _ptr(save),
                      _ptr(a_t), _ptr(c_t), _ptr(hn), _ptr(cn), _ptr(ws.sync) if fused else None)
    u = torch.empty(T, N, dtype=torch.float32, device=dev)
    q = torch.empty(KM, T, N, dtype=torch.float32, device=dev) if need_moments else None
    tens = _tensors(params)
    _timed_call("dlap_cell_fwd", dev, ws.pp, _ptr(x), _ptr(mask), _ptr(index), _ptr(ximg), C.byref(tens), C.byref(mio), _ptr(u),
                _ptr(q), _PRECISIONS[precision], float(dropout_p), int(dropout_seed), _ptr(dropout_seed_dev),
                _stream(dev))
    ws.launches += 1 if (fused or tables is not None) else 2
    return CellPass(u, q, save, a_t, c_t, hn, cn, index, mio, precision, ximg)


@dataclass
class Reduced:
    """Outputs of dlap_step_reduce: per-month statistics, portfolio returns, [L_c, L_u, Sharpe, N_bar, ...], the
    demeaned weights, and what the backward pass needs (unit loss coefficients G, the two dS sums)."""
    row_stats: torch.Tensor
    pr: torch.Tensor
    scalars: torch.Tensor
    w: torch.Tensor
    G: Optional[torch.Tensor]
    dS2: Optional[torch.Tensor]


def step_reduce(ws: Workspace, cp: CellPass, R, mask, weighted_loss: bool, want_backward: bool,
                shard: Optional["MonthShard"] = None, demean: bool = True, q_dense=None) -> Reduced:
    """dlap_step_reduce: every sum over stocks and over months, one launch (two around the all-reduce of the
    per-stock sums when the panel is a month shard).  q_dense: moment outputs [8,T,N] that did not come from the
    per-cell kernel (they replace cp.q)."""
    flags = (0 if demean else _lib.DLAP_REDUCE_NO_DEMEAN) | (_lib.DLAP_REDUCE_Q_DENSE if q_dense is not None else 0)
    q = cp.q if q_dense is None else q_dense
    stocks = shard if isinstance(shard, StockShard) else None
    if stocks is not None:
        flags |= _lib.DLAP_REDUCE_STOCK_SHARD
        shard = None
    dev = R.device
    T, N = R.shape
    row_stats = torch.empty(T, RS_W, dtype=torch.float32, device=dev)
    pr = torch.empty(T if shard is None else shard.T_total, dtype=torch.float32, device=dev)   # month shard: the whole series'
    xbuf = None if shard is None else torch.empty((KM + 2) * N + shard.T_total, dtype=torch.float32, device=dev)
    row_sums = torch.empty(T, 4, dtype=torch.float64, device=dev) if stocks is not None else None
    scalars = torch.empty(SC_W, dtype=torch.float32, device=dev)
    w = torch.empty(T, N, dtype=torch.float32, device=dev)
    e = torch.empty(KM + 1, N, dtype=torch.float32, device=dev)
    t_i = torch.empty(N, dtype=torch.float32, device=dev)
    G = torch.empty(N, GS, dtype=torch.float32, device=dev)
    dS2 = torch.empty(T, 2, dtype=torch.float64, device=dev) if want_backward else None
    if stocks is not None:          # [T,2] dS sums, then the two sums of e^2: one buffer, one all-reduce
        dS2 = torch.empty(2 * T + 2, dtype=torch.float64, device=dev)

    def launch(lo, hi):
        _timed_call("dlap_step_reduce", dev, ws.pp, _ptr(cp.u), _ptr(q), _ptr(R), _ptr(mask), _ptr(cp.index),
                    int(weighted_loss), flags, _ptr(shard.n_bar) if shard is not None else None, lo, hi, _ptr(row_stats),
                    _ptr(pr), _ptr(scalars), _ptr(w), _ptr(ws.col_part), _ptr(e), _ptr(t_i), _ptr(G),
                    _ptr(ws.loss_part), _ptr(dS2), _ptr(ws.row_part), _ptr(ws.bar),
                    C.byref(cp.mio) if shard is not None else None, _ptr(xbuf), _ptr(row_sums),
                    stocks.n_total if stocks is not None else 0,
                    C.byref(stocks.peers.c) if (stocks is not None and stocks.peers is not None) else None, _stream(dev))
        ws.launches += 1

    if stocks is not None and stocks.peers is not None:
        if T > stocks.peers.T_max:
            raise ValueError("the stock shard's peer buffers were made for fewer months")
        launch(0, 5)            # both exchanges inside the launch, over peer memory
    elif stocks is not None:
        launch(0, 0)
        with _timed_region("exchange_fwd", dev):
            torch.distributed.all_reduce(row_sums, group=stocks.group)
        launch(1, 4)
        with _timed_region("exchange_fwd2", dev):
            torch.distributed.all_reduce(dS2, group=stocks.group)
        launch(5, 5)
    elif shard is None:
        launch(0, 4)
    else:       # exchange 1 (the only one of the forward pass): per-stock sums over ALL months + the series' returns
        launch(0, 2)
        with _timed_region("exchange_fwd", dev):
            torch.distributed.all_reduce(xbuf, group=shard.group)
        launch(3, 4)
    return Reduced(row_stats, pr, scalars, w, G, dS2)


RS_SR = 6     # row_stats: sum_i m R of the month


def residual_terms(w, R, mask):
    """compute_residual_loss (src/model.py:435-483) from the weights of a step: the loss L, and per month c_t and
    beta_t with dL/dw_ti = c_t (R_ti - beta_t w_ti) on the listed cells.  Months with fewer than two listed stocks
    are left out; months whose weights vanish (w.w <= 1e-8) only count in the denominator.  Per-month sums in fp64
    ([T]-sized torch ops and four [T,N] reductions: this is a regulariser the reference's configurations leave at 0,
    not a hot path); no host synchronisation, so it can be captured in a CUDA graph."""
    m = mask.to(torch.float64)
    wd, Rd = w.to(torch.float64), R.to(torch.float64)
    n = m.sum(dim=1)
    ww = (wd * wd * m).sum(dim=1)
    Rw = (Rd * wd * m).sum(dim=1)
    RR = (Rd * Rd * m).sum(dim=1)
    zero, one = torch.zeros_like(n), torch.ones_like(n)
    B = n >= 2                                               # src/model.py:458-459
    A = B & (ww > 1e-8)                                      # :466
    ns = torch.where(B, n, one)
    beta = torch.where(A, Rw / torch.where(A, ww, one), zero)
    res = torch.where(A, (RR - Rw * beta) / ns, zero)        # mean (R - beta w)^2 over the month's listed stocks
    nA, nB = A.sum().to(torch.float64), B.sum().to(torch.float64)
    den = (torch.where(B, RR / ns, zero).sum() / nB.clamp(min=1.0)).clamp(min=1e-8)      # :478, :480
    loss = torch.where(nA > 0, res.sum() / nA.clamp(min=1.0) / den, torch.zeros_like(den))
    c = torch.where(A, -2.0 * beta / (nA.clamp(min=1.0) * den * ns), zero)
    return loss.to(torch.float32), c, beta


class GanStepFn(torch.autograd.Function):
    """(L_c, L_u, pr, weights, sharpe, h_n, c_n, L_res) = step(panel, params).  L_res: the residual-loss regulariser
    (src/model.py:435-483) when opts.residual, else a zero that carries no gradient.

    `cells`: a CellPass of the same panel and parameters made earlier (the evaluation cache of AssetPricingGAN), or
    None.  Three optional tensor inputs widen the fast path to the reference's other configurations:
      a_in [T,64], c_in [T,8] (or None)  the per-month biases come from the caller (a model without the LSTM,
                 src/model.py:200-205: a_t = W1[:,F:].macro_t + b1 by a plain matmul); `macro` is then None and the
                 parameter list holds the characteristics' columns of W1 / Wq only; their gradients are returned;
      q_in [8,T,N] dense                 the moment outputs come from the caller (a moment network with hidden layers,
                 src/model.py:119-127, evaluated with torch ops); its gradient is returned."""

    @staticmethod
    def forward(ctx, opts: StepOptions, cells: Optional[CellPass], macro, x, R, mask, h0, c0, a_in, c_in, q_in,
                *params):
        dev = x.device
        T, N, F = x.shape
        tables = a_in is not None
        M = 0 if (macro is None or tables) else macro.shape[1]
        H = 0 if M == 0 else params[1].shape[1]
        sh = opts.shard
        if (tables or q_in is not None) and sh is not None:
            raise NotImplementedError("shards with caller-made bias tables / moment outputs")
        if isinstance(sh, StockShard) and not opts.demean:
            raise NotImplementedError("stock shards with normalize_w=False")
        if opts.residual and (sh is not None or not opts.demean):
            raise NotImplementedError("residual_loss_factor > 0 with a sharded panel or normalize_w=False")
        want_backward = any(ctx.needs_input_grad)
        kernel_moments = opts.need_moments and q_in is None
        with torch.cuda.device(dev):
            ws = workspace(T, N, F, M, H, dev)
            cp = cells
            if cp is None:
                cp = cell_forward(ws, None if tables else macro, x, mask, params, h0, c0, opts.precision, kernel_moments,
                                  opts.dropout_p, opts.dropout_seed, opts.dropout_seed_dev, sh,
                                  tables=(a_in, c_in) if tables else None)
            q_dense = q_in.detach().contiguous() if (q_in is not None and opts.need_moments) else None
            red = step_reduce(ws, cp, R, mask, opts.weighted_loss, want_backward, sh, opts.demean, q_dense)
        ctx.set_materialize_grads(False)      # unused outputs come back as None, not as zero tensors (fill kernels)
        ctx.save_for_backward(x, R, mask, macro, h0, c0, *params)
        # What the backward pass reads -- and nothing that is an OUTPUT of this function: an output held by ctx is a
        # reference cycle (output -> grad_fn -> ctx -> output) that keeps the graph, and with it the parameters'
        # AccumulateGrad nodes and their stream, alive until the garbage collector runs.  u and q are dead here.
        ctx.opts, ctx.ws = opts, ws
        ctx.keep = (cp.save, cp.a_t, cp.c_t, cp.index, cp.mio, cp.q is not None, red.row_stats, red.scalars, red.G,
                    red.dS2, tables, q_dense is not None, cp.ximg)
        scalars = red.scalars
        l_c, l_u, sharpe = scalars[SC_LC], scalars[SC_LU], scalars[SC_SHARPE]
        if opts.residual:
            l_res, res_c, res_beta = residual_terms(red.w, R, mask)
            ctx.res = (res_c, res_beta, cp.u)              # [T] fp64 each; u: what dlap_cell_bwd needs for the term in u
            ctx.mark_non_differentiable(red.w, sharpe, cp.hn, cp.cn)
        else:
            l_res, ctx.res = scalars[7], None              # SC_ZERO: always 0.0
            ctx.mark_non_differentiable(red.w, sharpe, cp.hn, cp.cn, l_res)
        return l_c, l_u, red.pr, red.w, sharpe, cp.hn, cp.cn, l_res

    @staticmethod
    def backward(ctx, g_lc, g_lu, g_pr, *rest):
        x, R, mask, macro, h0, c0, *params = ctx.saved_tensors
        opts, ws = ctx.opts, ctx.ws
        g_lres = rest[4] if len(rest) > 4 else None
        save, a_t, c_t, index, mio, has_q, row_stats, scalars, G, dS2, tables, q_given, ximg = ctx.keep
        sh = opts.shard
        stocks = isinstance(sh, StockShard)
        # rows of the da_t | dc_t table that is exchanged: the whole series' (month shards) / this panel's (stock shards)
        T_rows = 0 if sh is None else (x.shape[0] if stocks else sh.T_total)
        dev = x.device
        T, N, F = x.shape
        M = ws.plan.dims.M
        NP = 11                                             # inputs in front of the parameters
        need_a, need_c, need_q = ctx.needs_input_grad[8:11]
        need = ctx.needs_input_grad[NP:]
        want_sdf = any(need[:10]) or need_a
        want_mom = (any(need[10:]) or need_c) and has_q
        if not (want_sdf or want_mom or (need_q and q_given)):
            return (None,) * (NP + len(params))
        prec = _PRECISIONS[_BWD_PRECISION[opts.precision]]

        def scalar(t):
            return None if t is None else t.detach().reshape(1).to(torch.float32).contiguous()

        g_lc, g_lu = scalar(g_lc), scalar(g_lu)
        if g_pr is not None:
            g_pr = g_pr.detach().to(torch.float32).contiguous()
        res_b = u_fwd = None
        if ctx.res is not None and g_lres is not None and want_sdf:
            # dL_res/dw_ti = c_t (R_ti - beta_t w_ti) on the listed cells and w = m (u - mu_t), so u receives
            # a_t (R_ti - Rbar_t) + b_t (u_ti - mu_t) with a_t = g c_t, b_t = -g c_t beta_t.  The first term has the form
            # of the portfolio return's (dP_t (R - Rbar_t), dP_t = dS_t N_bar / N_t when weighted): it joins g_pr.
            res_c, res_beta, u_fwd = ctx.res
            a = g_lres.detach().to(torch.float64) * res_c
            res_b = (-(a * res_beta)).to(torch.float32).contiguous()
            if opts.weighted_loss:
                a = a * (row_stats[:, RS_NT].to(torch.float64) / scalars[SC_NBAR].to(torch.float64))
            g_pr = a.to(torch.float32) if g_pr is None else g_pr + a.to(torch.float32)
        grad_a = grad_c = grad_q = None
        grads = [None] * len(params)
        with torch.cuda.device(dev):
            st = _stream(dev)
            if want_sdf or want_mom:
                tens = _tensors(params)
                _timed_call("dlap_cell_bwd", dev, ws.pp, _ptr(x), _ptr(R), _ptr(mask), _ptr(index), _ptr(ximg), C.byref(tens),
                            C.byref(mio), _ptr(row_stats), _ptr(dS2), _ptr(g_lc) if (has_q or q_given) else None,
                            _ptr(g_lu), _ptr(g_pr), _ptr(G), _ptr(scalars), int(opts.weighted_loss), int(want_sdf),
                            int(want_mom), _ptr(ws.wpart), _ptr(ws.dadc_part), prec, float(opts.dropout_p),
                            int(opts.dropout_seed), _ptr(opts.dropout_seed_dev), _ptr(res_b), _ptr(u_fwd), st)
                ws.launches += 1
                # All gradient tensors are views of one flat buffer.  Month-sharded: the buffer starts with the
                # [T_total,72] da_t | dc_t rows (this rank fills its months) and is zero-filled, so that exchange 2 is
                # ONE all-reduce over it -- gradients that are computed later (LSTM, b1, bq) are still zero.
                # Otherwise no fill: the kernels write every element of every gradient that is handed back.
                sizes = [0 if p is None else p.numel() for p in params]
                peer_bwd = stocks and sh.peers is not None and want_sdf and M > 0
                head = 0 if peer_bwd else T_rows * 72
                buf = (torch.empty if (sh is None or peer_bwd) else torch.zeros)(head + sum(sizes), dtype=torch.float32,
                                                                                device=dev)
                flat_g = buf[head:]
                grads, off = [], 0
                for p, n in zip(params, sizes):
                    grads.append(None if p is None else flat_g[off:off + n].view_as(p))
                    off += n
                gt = _tensors(grads)
                if tables:           # the month-level backward is the caller's (autograd of how it made a_t / c_t)
                    dadc = torch.empty(T, 72, dtype=torch.float32, device=dev)
                    _timed_call("dlap_grad_finish", dev, ws.pp, _ptr(ws.wpart), _ptr(ws.dadc_part), _ptr(index), prec,
                                int(want_sdf), int(want_mom), C.byref(gt), _ptr(dadc), 1, st)
                    ws.launches += 1
                    grad_a = dadc[:, :64] if (need_a and want_sdf) else None
                    grad_c = dadc[:, 64:] if (need_c and want_mom) else None
                elif sh is None or (stocks and sh.peers is not None and want_sdf and M > 0):
                    # one device -- or stock shards over peer memory: the exchange is inside the BPTT kernel
                    dadc = torch.empty(T, 72, dtype=torch.float32, device=dev)
                    _timed_call("dlap_macro_bwd", dev, ws.pp, _ptr(macro), C.byref(tens), _ptr(h0), _ptr(c0), _ptr(save),
                                _ptr(dadc), _ptr(ws.dpre), int(want_sdf), int(want_mom), C.byref(gt), _ptr(ws.wpart),
                                _ptr(ws.dadc_part), _ptr(index), prec, _ptr(ws.counter),
                                C.byref(sh.peers.g) if sh is not None else None, st)
                    # the month-summed gradients ride in the BPTT kernel when its grid is launched co-resident
                    ws.launches += 1 if (want_sdf and M > 0 and ws.plan.cooperative) else 2
                else:   # exchange 2: per-cell weight grads + every month's da_t | dc_t row
                    dadc_all = buf[:head].view(T_rows, 72)
                    t0, stride = (0, 1) if stocks else (sh.t0, sh.stride)
                    _timed_call("dlap_grad_finish", dev, ws.pp, _ptr(ws.wpart), _ptr(ws.dadc_part), _ptr(index), prec,
                                int(want_sdf), int(want_mom), C.byref(gt), _ptr(dadc_all[t0:]), int(stride), st)
                    if not opts.demean and want_sdf:
                        grads[9].copy_(_bias3_grad(opts, row_stats, scalars, dS2, g_lc if has_q or q_given else None, g_lu, g_pr))
                    wsm = workspace(T_rows, N, F, M, ws.plan.dims.H, dev)
                    with _timed_region("exchange_bwd", dev):
                        torch.distributed.all_reduce(buf, group=sh.group)
                    _timed_call("dlap_macro_bwd", dev, wsm.pp, _ptr(macro), C.byref(tens), _ptr(h0), _ptr(c0),
                                _ptr(save), _ptr(dadc_all), _ptr(wsm.dpre), int(want_sdf), int(want_mom), C.byref(gt),
                                None, None, None, prec, None, None, st)
                    ws.launches += 3 if (want_sdf and M > 0) else 2
                if not opts.demean and want_sdf and sh is None and grads[9] is not None:
                    grads[9].copy_(_bias3_grad(opts, row_stats, scalars, dS2, g_lc if has_q or q_given else None, g_lu, g_pr))
            if need_q and q_given and g_lc is not None:
                # dL/dq_kti = dL/dL_c * G_ki R_ti m_ti S_t  (autograd of src/model.py:425-431; G holds unit coefficients)
                rms = (R * mask) * row_stats[:, RS_S].unsqueeze(1)
                grad_q = (G[:, :KM].t().unsqueeze(1) * rms.unsqueeze(0)) * g_lc
        out = []
        for i, gr in enumerate(grads):
            live = need[i] and gr is not None and ((i < 10 and want_sdf) or (i >= 10 and want_mom))
            out.append(gr if live else None)
        return (None,) * 8 + (grad_a, grad_c, grad_q) + tuple(out)


def _bias3_grad(opts, row_stats, scalars, dS2, g_lc, g_lu, g_pr):
    """d loss / d output_proj.bias when the weights are not demeaned (normalize_w = False): every u shifts by db3, so
    it is sum_t dP_t sum_i m R -- a month-level sum, done with a handful of [T]-sized torch ops (with demeaning it is
    identically zero and the kernels write that)."""
    dS = torch.zeros(dS2.shape[0], dtype=torch.float64, device=dS2.device)
    if g_lc is not None:
        dS = dS + g_lc.double() * dS2[:, 0]
    if g_lu is not None:
        dS = dS + g_lu.double() * dS2[:, 1]
    if g_pr is not None:
        dS = dS + g_pr.double()
    if opts.weighted_loss:
        dS = dS * (scalars[SC_NBAR].double() / row_stats[:, RS_NT].double())
    return (dS * row_stats[:, RS_SR].double()).sum().to(torch.float32).reshape(1)


def sdf_weights(macro, x, mask, params, h0=None, c0=None, normalized=False, precision="fp32",
                with_moments=False, tables=None, demean=True):
    """Inference path of SDFNetwork.forward / get_weights (src/model.py:221-281, :565-594):
    weights [T,N] (demeaned unless demean=False, optionally |w|-normalised per month) and the per-cell pass they
    came from (with_moments: the moment outputs are computed too, for a forward() of the same panel that may follow)."""
    dev = x.device
    T, N, F = x.shape
    M = 0 if (macro is None or tables is not None) else macro.shape[1]
    H = 0 if M == 0 else params[1].shape[1]
    with torch.cuda.device(dev):
        ws = workspace(T, N, F, M, H, dev)
        cp = cell_forward(ws, None if tables is not None else macro, x, mask, params, h0, c0, precision, with_moments,
                          tables=tables)
        w = torch.empty(T, N, dtype=torch.float32, device=dev)
        _lib.call("dlap_weights", ws.pp, _ptr(cp.u), _ptr(mask), _ptr(cp.index), int(normalized), int(demean), _ptr(w),
                  _stream(dev))
        ws.launches += 1
    return w, cp
