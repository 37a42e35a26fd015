"""The epoch of the reference's training loop as one CUDA graph.

``train_epoch`` of ``src/train.py:45-103`` is, per epoch: ``.to(device)`` of the panel, ``zero_grad``, forward,
backward, ``clip_grad_norm_`` over the scope's parameters, ``Adam.step`` and five ``.item()`` reads.  With the step
itself at a few hundred microseconds those host round trips are the epoch.  Here

* ``FusedClipAdam`` -- ``torch.optim.Adam`` with its defaults (no weight decay, no amsgrad) and, optionally, the
  gradient clipping in front of it, as ONE kernel launch over all tensors (``dlap_clip_adam``); constructor,
  ``step`` / ``zero_grad`` / ``param_groups`` / ``state_dict`` follow ``torch.optim.Adam`` so that the reference's
  loops can hold one without knowing;
* ``train_epoch`` -- same signature and same returned dict as the reference's: the panel is uploaded once per panel
  (not once per epoch), zero_grad + forward + backward + clip + Adam are captured once per (panel, phase, scope,
  optimiser) and replayed, and the epoch's numbers come back in ONE device-to-host copy.
"""
from __future__ import annotations

import ctypes as C
import weakref
from typing import Dict, Iterable, Optional

import numpy as np
import torch

from . import _lib
from . import functional as Fn
from ._lib import ADAM_MAX_TENSORS, DlapAdamTensors


class FusedClipAdam:
    """Adam (torch.optim.Adam defaults; src/train.py:210-211 builds ``optim.Adam(params, lr=lr)``), the update of all
    tensors in one launch.  ``max_norm``: also do ``clip_grad_norm_(params, max_norm)`` (src/train.py:87-92) in the same
    launch, on the gradients of exactly these parameters; 0 = no clipping (the caller clips, as the reference's own
    train_epoch does).  Parameters whose ``.grad`` is None are skipped, like torch.  After a step ``grad_norm`` holds the
    total gradient norm (device tensor; before clipping, like clip_grad_norm_'s return value)."""

    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 max_norm: float = 0.0):
        self.params = list(params)
        if not self.params:
            raise ValueError("optimizer got an empty parameter list")
        if len(self.params) > ADAM_MAX_TENSORS:
            raise ValueError(f"at most {ADAM_MAX_TENSORS} tensors per FusedClipAdam")
        self.defaults = {"lr": lr, "betas": tuple(betas), "eps": eps, "max_norm": float(max_norm)}
        self.param_groups = [dict(self.defaults, params=self.params)]
        self.state: Dict[torch.nn.Parameter, Dict[str, torch.Tensor]] = {}
        self._step = None          # int64 [len(params)] on the device: updates so far, per tensor (as torch counts)
        self.grad_norm = None      # float32 [1] on the device
        self.launches = 0
        self._slot = {p: i for i, p in enumerate(self.params)}

    # -- state ------------------------------------------------------------------------------------------
    def ensure_state(self) -> None:
        """Allocate the moments (zero) and the step counter.  Must have happened before a stream capture: an
        allocation-with-fill inside the capture would be replayed, i.e. reset the moments every step."""
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise ValueError("FusedClipAdam needs contiguous float32 CUDA parameters")
            if p not in self.state:
                self.state[p] = {"exp_avg": torch.zeros_like(p, memory_format=torch.contiguous_format),
                                 "exp_avg_sq": torch.zeros_like(p, memory_format=torch.contiguous_format)}
        dev = self.params[0].device
        if self._step is None:
            self._step = torch.zeros(len(self.params), dtype=torch.int64, device=dev)
            self.grad_norm = torch.zeros(1, dtype=torch.float32, device=dev)

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.params:
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.detach_().zero_()

    @torch.no_grad()
    def step(self, max_norm: Optional[float] = None) -> None:
        self.ensure_state()
        g = self.param_groups[0]
        live = [p for p in self.params if p.grad is not None]
        if not live:
            return
        t = DlapAdamTensors()
        t.count = len(live)
        for k, p in enumerate(live):
            if not p.grad.is_contiguous() or p.grad.dtype != torch.float32:
                raise ValueError("FusedClipAdam needs contiguous float32 gradients")
            st = self.state[p]
            t.param[k], t.grad[k] = p.data_ptr(), p.grad.data_ptr()
            t.exp_avg[k], t.exp_avg_sq[k] = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
            t.step[k] = self._step.data_ptr() + 8 * self._slot[p]
            t.numel[k] = p.numel()
        dev = live[0].device
        with torch.cuda.device(dev):
            _lib.call("dlap_clip_adam", C.byref(t), float(g["max_norm"] if max_norm is None else max_norm),
                      float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
                      self.grad_norm.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
        bump_versions(live)
        self.launches += 1

    # -- torch.optim.Adam-shaped checkpoints -------------------------------------------------------------------
    def state_dict(self) -> dict:
        self.ensure_state()
        step = self._step.to(torch.float32).cpu()
        return {"state": {i: {"step": step[i].clone(), "exp_avg": self.state[p]["exp_avg"].clone(),
                              "exp_avg_sq": self.state[p]["exp_avg_sq"].clone()} for i, p in enumerate(self.params)},
                "param_groups": [{k: v for k, v in self.param_groups[0].items() if k != "params"}
                                 | {"params": list(range(len(self.params)))}]}

    def load_state_dict(self, sd: dict) -> None:
        self.ensure_state()
        for i, p in enumerate(self.params):
            s = sd["state"].get(i)
            if s is None:
                continue
            self.state[p]["exp_avg"].copy_(s["exp_avg"])
            self.state[p]["exp_avg_sq"].copy_(s["exp_avg_sq"])
            self._step[i] = int(float(s["step"]))
        for k in ("lr", "betas", "eps", "max_norm"):
            if k in sd["param_groups"][0]:
                self.param_groups[0][k] = sd["param_groups"][0][k]


def bump_versions(params) -> None:
    """A kernel wrote these parameters through raw pointers: bump their version counters, as an in-place torch op
    would have, so that everything keyed on them -- autograd's saved-tensor checks, the model's evaluation cache --
    sees the change."""
    params = tuple(params)
    torch._C._autograd._unsafe_set_version_counter(params, tuple(p._version + 1 for p in params))


# ---- train_epoch ---------------------------------------------------------------------------------------------
_resident = {}      # id(host tensor) -> (weakref to it, its version, device copy): panels uploaded once, not per epoch


def _on_device(t: Optional[torch.Tensor], device: torch.device) -> Optional[torch.Tensor]:
    """`t.to(device)` of src/train.py:69-75, done once per host tensor (and again only when it changed)."""
    if t is None or t.device == device:
        return t
    hit = _resident.get(id(t))
    if hit is not None and hit[0]() is t and hit[1] == t._version and hit[2].device == device:
        return hit[2]
    d = t.to(device)
    _resident[id(t)] = (weakref.ref(t), t._version, d)
    for k in [k for k, v in _resident.items() if v[0]() is None]:
        del _resident[k]
    return d


def _scope_params(model, scope: str):
    if scope == 'sdf':
        return list(model.sdf_net.parameters())
    if scope == 'moment':
        return list(model.moment_net.parameters())
    return list(model.parameters())


def train_epoch(model, optimizer, data: dict, device: torch.device, phase: str = 'conditional',
                grad_clip: float = 1.0, scope: str = 'all') -> dict:
    """Drop-in for ``train_epoch`` of src/train.py:45-103 (same arguments, same returned dict) for a
    ``FusedClipAdam`` optimiser: one CUDA-graph replay and one device-to-host copy per epoch.

    The graph (zero_grad, forward of ``phase``, backward, clipping over the scope's parameters, Adam) is captured on
    the first call for a given (panel, phase, scope, optimiser, set of trainable parameters) and kept on the model.
    The clipping scope must be the optimiser's own parameters -- which is how train_3phase and the notebooks call it
    (sdf optimiser with scope 'sdf', moment optimiser with scope 'moment', one optimiser over everything with 'all')."""
    if not isinstance(optimizer, FusedClipAdam):
        raise TypeError("this train_epoch drives a FusedClipAdam (construct it where the reference constructs optim.Adam)")
    device = torch.device(device)
    model.train()
    macro = _on_device(data.get('macro_features'), device)
    x = _on_device(data['individual_features'], device)
    R = _on_device(data['returns'], device)
    mask = _on_device(data['mask'], device)
    scope_set = {id(p) for p in _scope_params(model, scope) if p.requires_grad}
    opt_set = {id(p) for p in optimizer.params if p.requires_grad}
    if scope_set != opt_set:
        raise NotImplementedError("train_epoch fuses clip_grad_norm_ with the optimiser step: the clipping scope must "
                                  "be exactly the optimiser's trainable parameters")
    trainable = tuple(p.requires_grad for p in model.parameters())
    key = (id(macro), id(x), id(R), id(mask), phase, scope, id(optimizer), trainable, float(grad_clip))
    graphs = model.__dict__.setdefault("_epoch_graphs", {})
    hit = graphs.get(key)
    if hit is None or any(r() is not t for r, t in zip(hit[1], (macro, x, R, mask)) if t is not None):
        gs = model.graphed_step(macro, x, R, mask, phase=phase, optimizer=optimizer, max_norm=float(grad_clip))
        pinned = torch.empty(gs.report.numel(), dtype=torch.float32).pin_memory()
        hit = (gs, [None if t is None else weakref.ref(t) for t in (macro, x, R, mask)], pinned)
        if len(graphs) >= 8:
            graphs.pop(next(iter(graphs)))
        graphs[key] = hit
    gs, _, pinned = hit
    gs.step()
    pinned.copy_(gs.report, non_blocking=True)        # the graph packed the epoch's numbers (GraphedStep.report)
    torch.cuda.current_stream(device).synchronize()
    vals = pinned.numpy()
    pr = vals[5:].astype(np.float64)
    std = pr.std(ddof=1) if pr.size > 1 else float('nan')
    sharpe = 0.0 if std < 1e-8 else float(pr.mean() / std)          # compute_sharpe, src/train.py:29-34
    return {'loss': float(vals[0]), 'loss_unc': float(vals[1]), 'loss_cond': float(vals[2]), 'loss_residual': float(vals[4]),
            'sharpe': sharpe, 'grad_norm': float(vals[3])}
