"""Drop-in mirror of the reference's model classes (``src/model.py``) on the B200 kernels.

Same constructors, same attribute names, same ``state_dict`` keys/shapes and the same default
initialisation (real ``nn.LSTM`` / ``nn.Linear`` modules hold the parameters and are created in
the reference's order, so ``torch.manual_seed(s)`` gives identical initial weights).  The math
does not run through those modules: every forward goes through ``libdlap_b200.so``.

Configurations outside the fast path raise ``NotImplementedError`` rather
than silently taking another route; there is no CPU path.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from . import functional as Fn


def _unsupported(what: str):
    raise NotImplementedError(
        f"{what} is outside the B200 fast path of this build (see DESIGN.md, 'out of scope'); "
        "there is no fallback path.")


def _hidden_pair(hidden, H, dev):
    if hidden is None:
        return None, None
    h0, c0 = hidden
    return Fn._flat_state(h0, H, dev), Fn._flat_state(c0, H, dev)


class MacroLSTM(nn.Module):
    """src/model.py:21-84."""

    def __init__(self, input_dim: int, hidden_dims: List[int], dropout: float = 0.0):
        super().__init__()
        self.hidden_dims = hidden_dims
        self.num_layers = len(hidden_dims)
        self.lstm = nn.LSTM(input_size=input_dim, hidden_size=hidden_dims[-1],
                            num_layers=self.num_layers, batch_first=True,
                            dropout=dropout if self.num_layers > 1 else 0)
        self.output_dim = hidden_dims[-1]

    def _params(self):
        if self.num_layers != 1:
            _unsupported("a multi-layer macro LSTM")
        if self.output_dim > 8:
            _unsupported("an LSTM wider than 8 units")
        l = self.lstm
        return [l.weight_ih_l0, l.weight_hh_l0, l.bias_ih_l0, l.bias_hh_l0]

    def forward(self, x: torch.Tensor, hidden=None):
        """x [T, input_dim] -> (output [T, H], (h_n, c_n) each [1,1,H]).  Inference only when
        called on its own (the training gradient flows through AssetPricingGAN.forward)."""
        if not x.is_cuda or x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 2:
            raise ValueError("macro features must be a contiguous float32 CUDA [T,M] tensor")
        H = self.output_dim
        lstm = [p.detach() for p in self._params()]
        dev = x.device
        T, M = x.shape
        # stand-alone call: only the LSTM part of the macro kernel matters; feed neutral MLP params
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        params = lstm + [z(64, 1 + H), z(64), None, None, None, None, z(8, M + 1), z(8)]
        h0, c0 = _hidden_pair(hidden, H, dev)
        with torch.cuda.device(dev):
            ws = Fn.workspace(T, 1, 1, M, H, dev)
            save, _, _, hn, cn = Fn.macro_forward(ws, x, params, h0, c0)
        return save[:, :H], (hn.view(1, 1, H), cn.view(1, 1, H))

    def init_hidden(self, device: torch.device):
        h0 = torch.zeros(self.num_layers, 1, self.hidden_dims[-1], device=device)
        return (h0, torch.zeros_like(h0))


class MomentNetwork(nn.Module):
    """src/model.py:87-161.  Stand-alone calls on an arbitrary [T,N,D] tensor use torch ops
    (this module is not on the step's hot path: AssetPricingGAN.forward fuses the moment net into
    the per-cell kernel and never materialises its [T,N,M+F] input)."""

    def __init__(self, input_dim: int, hidden_dims: List[int], num_moments: int,
                 use_rnn: bool = False, rnn_hidden_dims: List[int] = None, dropout: float = 0.05):
        super().__init__()
        self.use_rnn = use_rnn
        self.num_moments = num_moments
        layers = []
        prev = input_dim
        for hd in hidden_dims:
            layers += [nn.Linear(prev, hd), nn.ReLU(), nn.Dropout(dropout)]
            prev = hd
        self.fc_layers = nn.Sequential(*layers) if layers else nn.Identity()
        self.output_proj = nn.Linear(prev if hidden_dims else input_dim, num_moments)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        T, N, D = x.shape
        h = x.view(T * N, D)
        if not isinstance(self.fc_layers, nn.Identity):
            h = self.fc_layers(h)
        return torch.tanh(self.output_proj(h)).view(T, N, -1).permute(2, 0, 1)

    def moments_by_month_bias(self, macro: Optional[torch.Tensor], x: torch.Tensor) -> torch.Tensor:
        """forward(cat([macro tiled over the stocks, x])) of src/model.py:514-521 for a network WITH hidden layers,
        without the [T,N,M+F] concatenation: the first Linear is split into its characteristics' columns (a GEMM
        over the cells) and its macro columns (a per-month bias).  Library GEMMs and elementwise torch ops: this
        variant of the moment network is not fused into the per-cell kernels.  -> contiguous [K,T,N]."""
        first, rest = self.fc_layers[0], self.fc_layers[1:]
        M = 0 if macro is None else macro.shape[1]
        if first.in_features != M + x.shape[2]:
            raise ValueError("the moment network's input width does not match the panel")
        h = x @ first.weight[:, M:].t()
        h = h + (first.bias if macro is None else torch.addmm(first.bias, macro, first.weight[:, :M].t())[:, None, :])
        h = rest(h)
        return torch.tanh(self.output_proj(h)).permute(2, 0, 1).contiguous()


class SDFNetwork(nn.Module):
    """src/model.py:164-281."""

    def __init__(self, macro_dim: int, individual_dim: int, hidden_dims: List[int],
                 use_rnn: bool = True, rnn_hidden_dims: List[int] = None, dropout: float = 0.05,
                 normalize_weights: bool = True):
        super().__init__()
        self.use_rnn = use_rnn
        self.normalize_weights = normalize_weights
        self.macro_dim = macro_dim
        self.individual_dim = individual_dim
        self.dropout_p = float(dropout)
        if use_rnn and rnn_hidden_dims and macro_dim > 0:
            self.macro_lstm = MacroLSTM(input_dim=macro_dim, hidden_dims=rnn_hidden_dims, dropout=dropout)
            processed = self.macro_lstm.output_dim
        else:
            self.macro_lstm = None
            processed = macro_dim
        layers = []
        prev = processed + individual_dim
        for hd in hidden_dims:
            layers += [nn.Linear(prev, hd), nn.ReLU(), nn.Dropout(dropout)]
            prev = hd
        self.fc_layers = nn.Sequential(*layers)
        self.output_proj = nn.Linear(prev, 1)
        self._hidden_dims = list(hidden_dims)

    # -- fast-path plumbing -------------------------------------------------------------------
    def _check(self):
        if len(self._hidden_dims) != 2 or max(self._hidden_dims) > 64 or min(self._hidden_dims) < 1:
            _unsupported(f"hidden_dim={self._hidden_dims} (two hidden layers of at most 64 units)")
        if self.individual_dim > 64:
            _unsupported("more than 64 characteristics")

    def mlp_tensors(self):
        """(W1 [64, in], b1 [64], W2 [64,64], b2 [64], W3 [1,64], b3 [1]): the MLP's parameters as the kernels want
        them.  Hidden layers narrower than 64 units (the notebooks' [64, 32]) are padded with zero units by
        differentiable torch ops: a unit with zero weights and zero bias outputs relu(0) = 0, passes nothing on
        (zero outgoing weights) and receives zero gradient, so the padded network IS the narrow one, and autograd
        cuts the kernels' gradients back to the parameters' shapes."""
        self._check()
        d1, d2 = self._hidden_dims
        fc0, fc3, out = self.fc_layers[0], self.fc_layers[3], self.output_proj
        pad = torch.nn.functional.pad
        w1 = fc0.weight if d1 == 64 else pad(fc0.weight, (0, 0, 0, 64 - d1))
        b1 = fc0.bias if d1 == 64 else pad(fc0.bias, (0, 64 - d1))
        w2 = fc3.weight if (d1, d2) == (64, 64) else pad(fc3.weight, (0, 64 - d1, 0, 64 - d2))
        b2 = fc3.bias if d2 == 64 else pad(fc3.bias, (0, 64 - d2))
        w3 = out.weight if d2 == 64 else pad(out.weight, (0, 64 - d2))
        return w1, b1, w2, b2, w3, out.bias

    def forward(self, macro_features, individual_features, mask, hidden=None):
        """-> (weights [T,N], new_hidden).  Stand-alone use is the inference path (no autograd
        graph is attached to the returned weights)."""
        return _weights(self, None, macro_features, individual_features, mask, hidden, False)[:2]


class KernelInputs:
    """How one configuration of the reference's model reaches the kernels (GanStepFn's arguments)."""
    __slots__ = ("macro", "a_in", "c_in", "q_in", "params", "H")


def _kernel_inputs(sdf: SDFNetwork, moment: Optional[MomentNetwork], macro, x, want_q: bool) -> KernelInputs:
    """moment None: the moment network takes no part (weights-only inference); the kernel ABI still gets neutral
    tensors where it wants some.  want_q: the moment outputs are needed (phase != 'unconditional')."""
    k = KernelInputs()
    dev = x.device
    F = x.shape[2]
    w1, b1, w2, b2, w3, b3 = sdf.mlp_tensors()
    M = macro.shape[1] if macro is not None else 0
    lstm_mode = macro is not None and sdf.macro_lstm is not None
    tables_mode = macro is not None and sdf.macro_lstm is None          # use_rnn=False, src/model.py:200-205
    exp_in = F + (sdf.macro_lstm.output_dim if lstm_mode else M)
    if sdf.fc_layers[0].in_features != exp_in:
        raise ValueError(f"the SDF network expects {sdf.fc_layers[0].in_features} inputs per stock-month, the panel "
                         f"provides {exp_in} (macro_features=None needs a model built with macro_feature_dim=0)")
    k.macro, k.a_in, k.c_in, k.q_in = (macro if lstm_mode else None), None, None, None
    k.H = sdf.macro_lstm.output_dim if lstm_mode else 0
    lstm = sdf.macro_lstm._params() if lstm_mode else [None] * 4
    if tables_mode:
        # x first, then the macro series (src/model.py:253-255): a_t = W1[:, F:] macro_t + b1 by one small matmul
        k.a_in = torch.addmm(b1, macro, w1[:, F:].t())
        w1 = w1[:, :F].contiguous()
        b1 = None
    mom_hidden = moment is not None and not isinstance(moment.fc_layers, nn.Identity)
    wq = bq = None
    if moment is not None and moment.num_moments != 8:
        _unsupported("num_condition_moment != 8")
    if moment is not None and not mom_hidden:
        mp = moment.output_proj
        if mp.in_features != M + F:
            raise ValueError("the moment network's input width does not match the panel")
        if tables_mode:          # macro first, then x (src/model.py:514-516)
            k.c_in = torch.addmm(mp.bias, macro, mp.weight[:, :M].t())
            wq = mp.weight[:, M:].contiguous()
        else:
            wq, bq = mp.weight, mp.bias
    elif mom_hidden and want_q:
        k.q_in = moment.moments_by_month_bias(macro, x)
    if wq is None and not tables_mode:
        # the macro kernel fills c_t whatever happens: neutral moment parameters
        wq = torch.zeros(8, M + F, dtype=torch.float32, device=dev)
        bq = torch.zeros(8, dtype=torch.float32, device=dev)
    k.params = lstm + [w1, b1, w2, b2, w3, b3, wq, bq]
    return k


def _weights(sdf: SDFNetwork, moment, macro, x, mask, hidden, normalized, precision="fp32"):
    """-> (weights, new_hidden, CellPass).  moment: the moment network, to have its outputs computed in the same
    per-cell pass (for a forward() on the same panel that may follow), or None."""
    Fn.check_panel(macro, x, None, mask)
    with torch.no_grad():
        k = _kernel_inputs(sdf, moment, macro, x, want_q=False)
    params = [p.detach() if p is not None else None for p in k.params]
    dev = x.device
    h0, c0 = _hidden_pair(hidden, k.H, dev) if k.H else (None, None)
    with_moments = moment is not None and isinstance(moment.fc_layers, nn.Identity)
    tables = None if k.a_in is None else (k.a_in, k.c_in)
    w, cp = Fn.sdf_weights(macro, x, mask, params, h0, c0, normalized, precision, with_moments=with_moments,
                           tables=tables, demean=sdf.normalize_weights)
    new_hidden = (cp.hn.view(1, 1, k.H), cp.cn.view(1, 1, k.H)) if k.H else None
    return w, new_hidden, cp


class StepOutputs(dict):
    """Forward dict of AssetPricingGAN (src/model.py:553-563).  ``moments`` [K,T,N] is produced on
    first access: no caller in the reference reads it (src/train.py:96-103, :144-153) and writing
    it every step would cost 32 B per stock-month."""

    def __init__(self, *a, moments_fn=None, **kw):
        super().__init__(*a, **kw)
        self._moments_fn = moments_fn
        dict.__setitem__(self, "moments", None)

    def __getitem__(self, k):
        v = dict.__getitem__(self, k)
        if k == "moments" and v is None and self._moments_fn is not None:
            v = self._moments_fn()
            dict.__setitem__(self, k, v)
        return v

    def get(self, k, default=None):
        return self[k] if k in self else default


class AssetPricingGAN(nn.Module):
    """src/model.py:284-617 (compute_residual_loss, :435-483, as functional.residual_terms + a term of the per-cell
    backward kernel)."""

    def __init__(self, config: Dict):
        super().__init__()
        self.config = config
        rnn_hidden = config.get('num_units_rnn', [4])
        if isinstance(rnn_hidden, int):
            rnn_hidden = [rnn_hidden]
        rnn_hidden_moment = config.get('num_units_rnn_moment', [32])
        if isinstance(rnn_hidden_moment, int):
            rnn_hidden_moment = [rnn_hidden_moment]
        dropout = config.get('dropout', 0.05)
        self.sdf_net = SDFNetwork(
            macro_dim=config.get('macro_feature_dim', 0),
            individual_dim=config['individual_feature_dim'],
            hidden_dims=config.get('hidden_dim', [64, 64]),
            use_rnn=config.get('use_rnn', True),
            rnn_hidden_dims=rnn_hidden if config.get('use_rnn', True) else None,
            dropout=dropout,
            normalize_weights=config.get('normalize_w', True))
        self._hidden_dim_moment = list(config.get('hidden_dim_moment', []))
        self.moment_net = MomentNetwork(
            input_dim=config.get('macro_feature_dim', 0) + config['individual_feature_dim'],
            hidden_dims=self._hidden_dim_moment,
            num_moments=config.get('num_condition_moment', 8),
            use_rnn=config.get('use_rnn_moment', True),
            rnn_hidden_dims=rnn_hidden_moment if config.get('use_rnn_moment', True) else None,
            dropout=dropout)
        self.residual_loss_factor = config.get('residual_loss_factor', 0.0)
        self.weighted_loss = config.get('weighted_loss', True)
        # kernel arithmetic of the per-cell contractions: "bf16x3" = tensor cores on hi/lo-split operands,
        # "fp32" = exact FFMA on CUDA cores, "auto" (default: the reference's own config dict carries no such
        # key) = bf16x3 whenever the tensor-core kernels take the shape (F <= 46), fp32 otherwise
        self.precision = config.get('b200_precision', 'auto')
        if self.precision not in ('auto', 'fp32', 'bf16x3'):
            raise ValueError(f"b200_precision must be 'auto', 'fp32' or 'bf16x3', not {self.precision!r}")
        self.month_shard: Optional[Fn.MonthShard] = None

    # -- fast-path plumbing -------------------------------------------------------------------
    def _zero(self, dev):
        z = getattr(self, "_zero_cache", None)
        if z is None or z.device != dev:
            z = self._zero_cache = torch.zeros((), dtype=torch.float32, device=dev)
        return z

    def _precision(self, F: int) -> str:
        if self.precision != 'auto':
            return self.precision
        return 'bf16x3' if F <= Fn.TC_MAX_F else 'fp32'

    def _inputs(self, macro, x, want_q: bool) -> "KernelInputs":
        return _kernel_inputs(self.sdf_net, self.moment_net, macro, x, want_q)

    def forward(self, macro_features, individual_features, returns, mask, hidden=None,
                phase: str = 'conditional'):
        """src/model.py:485-563.  Differentiable outputs: ``loss``, the two loss terms and ``portfolio_returns``;
        ``weights``, ``sharpe`` and ``hidden`` carry no autograd history (nothing in the reference's loops
        differentiates through them; a loss term built on them would get no gradient)."""
        Fn.check_panel(macro_features, individual_features, returns, mask)
        x = individual_features
        dev = x.device
        k = self._inputs(macro_features, x, want_q=(phase != 'unconditional'))
        H = k.H
        with_macro = H > 0
        h0, c0 = _hidden_pair(hidden, H, dev) if with_macro else (None, None)
        p_drop = self.sdf_net.dropout_p if self.training else 0.0
        seed_dev = getattr(self, "_seed_dev", None) if p_drop > 0 else None   # GraphedStep: seed in device memory
        seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if p_drop > 0 and seed_dev is None else 0   # CPU generator: no sync
        opts = Fn.StepOptions(weighted_loss=bool(self.weighted_loss),
                              need_moments=(phase != 'unconditional'), precision=self._precision(x.shape[2]),
                              dropout_p=p_drop, dropout_seed=seed, dropout_seed_dev=seed_dev,
                              shard=self.month_shard, demean=bool(self.sdf_net.normalize_weights),
                              residual=self.residual_loss_factor > 0)
        cells = self._cached_cells(macro_features, x, mask, hidden, opts)
        l_c, l_u, pr, w, sharpe, hn, cn, l_res = Fn.GanStepFn.apply(
            opts, cells, k.macro, x, returns, mask, h0, c0, k.a_in, k.c_in, k.q_in, *k.params)
        zero = self._zero(dev)
        if phase == 'unconditional':            # src/model.py:524-528
            loss_unc, loss_cond, total = l_u, zero, l_u
        elif phase == 'moment':                 # src/model.py:530-535
            loss_unc, loss_cond, total = zero, l_c, -l_c
        else:                                   # src/model.py:537-541
            loss_unc, loss_cond, total = l_u, l_c, l_c
        if self.residual_loss_factor > 0:       # src/model.py:543-548
            total = total + self.residual_loss_factor * l_res
        else:
            l_res = zero
        new_hidden = (hn.view(1, 1, H), cn.view(1, 1, H)) if with_macro else None

        def moments():
            if k.q_in is not None:
                return k.q_in
            mp = self.moment_net.output_proj
            M = macro_features.shape[1] if macro_features is not None else 0
            pre = x @ mp.weight[:, M:].t()
            bias = mp.bias if M == 0 else macro_features @ mp.weight[:, :M].t() + mp.bias
            pre = pre + (bias[:, None, :] if M else bias)
            return torch.tanh(pre).permute(2, 0, 1)

        return StepOutputs({
            'weights': w, 'loss': total, 'loss_unconditional': loss_unc, 'loss_conditional': loss_cond,
            'loss_residual': l_res, 'sharpe': sharpe, 'portfolio_returns': pr, 'hidden': new_hidden,
        }, moments_fn=moments)

    def graphed_step(self, macro_features, individual_features, returns, mask, hidden=None,
                     phase: str = 'conditional', optimizer=None, max_norm: Optional[float] = None) -> "GraphedStep":
        """CUDA-graph capture of zero_grad + forward + backward (+ clip + Adam) on a resident panel (see GraphedStep)."""
        return GraphedStep(self, macro_features, individual_features, returns, mask, hidden, phase,
                           optimizer=optimizer, max_norm=max_norm)

    # -- evaluation cache ---------------------------------------------------------
    # evaluate() of src/train.py:106-153 calls get_weights(normalized=True) and then forward(phase='conditional') on
    # the same panel with the same parameters, i.e. two full passes over every stock-month.  Under no_grad in eval
    # mode, get_weights therefore also computes the moment outputs in its per-cell pass and keeps (u, q, LSTM state)
    # keyed on the identity and version counters of panel and parameters; the forward() that follows finds them and
    # runs only the reductions.  Anything else -- another panel, an optimiser step, train mode, grad mode -- misses.
    eval_cache = True

    def _cache_key(self, macro, x, mask, hidden, precision):
        """(scalars, tensors): version counters and data pointers of panel and parameters, and the tensor objects
        themselves -- the cache keeps weak references to those and matches by identity, so a recycled id() or
        address of a dead tensor can never hit."""
        tensors = [t for t in (macro, x, mask) + (tuple(hidden) if hidden is not None else ())
                   + tuple(self.parameters()) if t is not None]
        return (tuple((t._version, t.data_ptr()) for t in tensors), macro is None, hidden is None, precision), tensors

    def _cached_cells(self, macro, x, mask, hidden, opts):
        hit = getattr(self, "_cells", None)
        self._cells = None                        # one use: the cache never outlives the evaluate() it was made for
        if hit is None or torch.is_grad_enabled() or self.training or opts.shard is not None:
            return None
        scalars, refs, cp = hit
        now, tensors = self._cache_key(macro, x, mask, hidden, opts.precision)
        if now != scalars or len(refs) != len(tensors) or any(r() is not t for r, t in zip(refs, tensors)):
            return None
        if opts.need_moments and cp.q is None:
            return None
        return cp

    def get_weights(self, macro_features, individual_features, mask, hidden=None, normalized: bool = False):
        """src/model.py:565-594."""
        precision = self._precision(individual_features.shape[2])
        keep = (self.eval_cache and not torch.is_grad_enabled() and not self.training and self.month_shard is None
                and not self._hidden_dim_moment and self.residual_loss_factor == 0)
        w, new_hidden, cp = _weights(self.sdf_net, self.moment_net if keep else None, macro_features,
                                     individual_features, mask, hidden, normalized, precision)
        self._cells = None
        if keep:
            import weakref
            scalars, tensors = self._cache_key(macro_features, individual_features, mask, hidden, precision)
            self._cells = (scalars, [weakref.ref(t) for t in tensors], cp)
        return w, new_hidden

    def get_sdf_factor(self, macro_features, individual_features, returns, mask, hidden=None,
                       normalized: bool = True):
        """src/model.py:596-617."""
        w, _ = self.get_weights(macro_features, individual_features, mask, hidden, normalized=normalized)
        return (w * returns * mask.float()).sum(dim=1)


class GraphedStep:
    """One training step -- ``model.zero_grad(); out = model(panel, phase); out['loss'].backward()``
    (src/train.py:77-84) -- captured once in a CUDA graph and replayed: the ~15 kernel launches of a step cost
    about as much host time as they run on the device, so a loop over a resident panel is otherwise bound by
    Python.  ``step()`` replays the graph and returns the forward dict (static tensors, overwritten by the next
    replay); ``p.grad`` of every parameter is overwritten, not accumulated.

    Drop every reference to the outputs of earlier eager steps (``out``, ``loss``) before constructing one: an
    autograd graph that is still alive keeps the parameters' AccumulateGrad nodes on the stream it ran on, and the
    capture then fails with a stream-dependency error.

    The panel tensors are the ones passed in (full-batch training on a resident panel, as the reference's
    loops do); copy new VALUES into them with ``.copy_()`` if needed (``step()`` notices a new version of the
    characteristics and rebuilds their X images in place), but keep the mask: its valid-cell index is part of the capture.  With dropout on (model.train(), p > 0) the seed lives in device memory and the graph
    itself advances it, so every replay draws a fresh mask; month-sharded ranks start from the same seed.  A
    month-sharded model captures its NCCL all-reduces with the kernels; every rank must then replay in lock-step.

    ``optimizer`` (a training.FusedClipAdam): the graph ends with ``clip_grad_norm_(its parameters, max_norm)`` +
    ``optimizer.step()`` (src/train.py:87-94) as one more launch, i.e. a replay is a whole training epoch of the
    reference's full-batch loop; ``optimizer.grad_norm`` then holds the gradient norm of the last replay.  The warm-up
    steps before the capture do not touch the parameters or the optimiser state."""

    def __init__(self, model: "AssetPricingGAN", macro_features, individual_features, returns, mask,
                 hidden=None, phase: str = 'conditional', warmup: int = 3, optimizer=None,
                 max_norm: Optional[float] = None):
        self.model, self.args = model, (macro_features, individual_features, returns, mask)
        self.seed = None
        self._one = None
        self.optimizer, self.max_norm = optimizer, max_norm
        if optimizer is not None:
            optimizer.ensure_state()          # allocations with a fill must not be part of the capture
        if model.training and model.sdf_net.dropout_p > 0:
            s0 = int(torch.randint(0, 2 ** 62, (1,)).item())
            if model.month_shard is not None:             # the same masks on every rank
                import torch.distributed as dist
                t = torch.tensor([s0], dtype=torch.int64, device=individual_features.device)
                dist.broadcast(t, src=dist.get_global_rank(model.month_shard.group, 0) if model.month_shard.group
                               is not None else 0, group=model.month_shard.group)
                s0 = int(t.item())
            self.seed = torch.tensor([s0], dtype=torch.int64, device=individual_features.device)
        self.kw = {"hidden": hidden, "phase": phase}
        dev = individual_features.device
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        Fn.force_x_image = True                            # this panel is stepped on many times: X images at once
        try:
            with torch.cuda.stream(side):                  # warm-up off the capture stream (allocator, index, images)
                for _ in range(max(warmup, 1)):
                    self._eager()
        finally:
            Fn.force_x_image = False
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        ws_launches = lambda: sum(w.launches for w in Fn._workspaces.values())
        self.graph = torch.cuda.CUDAGraph()
        model.zero_grad(set_to_none=True)
        l0 = ws_launches()
        try:        # the parameters' AccumulateGrad nodes were created on the warm-up stream: harmless here
            torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch(False)
        except AttributeError:
            pass
        o0 = optimizer.launches if optimizer is not None else 0
        with torch.cuda.graph(self.graph):
            self.out = self._eager(zero=False, apply=True)
        # kernels of this library inside one replay
        self.launches = ws_launches() - l0 + ((optimizer.launches - o0) if optimizer is not None else 0)
        # the capture holds the address of the panel's X images: new values copied into x need them rebuilt in place
        self._ximg = Fn.cached_panel_image(individual_features)
        self._x_version = individual_features._version

    def _eager(self, zero=True, apply=False):
        if zero:
            self.model.zero_grad(set_to_none=True)
        self.model._seed_dev = self.seed
        try:
            if self.seed is not None:
                self.seed.add_(0x9E3779B97F4A7C15 >> 2)      # next mask; part of the captured graph
            out = self.model(*self.args, **self.kw)
            if self._one is None:                       # root gradient: made once, not filled on every replay
                self._one = torch.ones_like(out['loss'])
            out['loss'].backward(gradient=self._one)
            if apply and self.optimizer is not None:
                self.optimizer.step(self.max_norm)
                # what an epoch reports (src/train.py:96-103), packed for ONE device-to-host copy:
                # loss, L_u, L_c, gradient norm, residual loss, then the T portfolio returns
                with torch.no_grad():
                    self.report = torch.cat([torch.stack([out['loss'], out['loss_unconditional'],
                                                          out['loss_conditional'], self.optimizer.grad_norm[0],
                                                          out['loss_residual']]),
                                             out['portfolio_returns']])
        finally:
            self.model._seed_dev = None
        return out

    def step(self):
        x = self.args[1]
        if self._ximg is not None and x._version != self._x_version:
            Fn.rebuild_panel_image(x, *self._ximg)
            self._x_version = x._version
        self.graph.replay()
        if self.optimizer is not None:      # the replay moved the parameters behind torch's back
            from .training import bump_versions
            bump_versions([p for p in self.optimizer.params if p.grad is not None])
            self.model._cells = None
        return self.out

    __call__ = step
