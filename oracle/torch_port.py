"""Torch-CPU restatement of the reference's adversarial-SDF step.  TEST INFRASTRUCTURE ONLY
(see ``oracle/__init__.py`` for who may import this).

Parity status: PINNED -- ``tests/test_oracle_golden.py`` checks every output of this file
against fixtures produced by the reference itself (``tests/golden/make_golden.py``): equal bit
for bit on the host that wrote them, within 4e-6 of each tensor's largest entry on a host whose
CPU makes torch pick other GEMM/LSTM kernels.

The reference's arithmetic for this path lives in PyTorch (``requirements.txt:5``
``torch>=1.9.0``, unpinned; this image has 2.11.0+cu128), so the restatement issues the same
torch ops in the same order and reproduces it to the last bit on CPU.  The code is functional
(a flat ``params`` dict keyed by the reference's ``state_dict`` names) rather than a module
tree.  Dropout is not restated: every parity comparison runs with dropout disabled
(BASELINE.json north_star), i.e. the reference in ``eval()`` mode or with ``dropout=0.0``.

Citations are ``/root/reference`` paths.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

# state_dict keys of the default configuration, in registration order (src/model.py:39-45,
# 208-219, 127-130).  fc_layers indices 0 and 3 come from the Linear/ReLU/Dropout interleave.
K_WIH = "sdf_net.macro_lstm.lstm.weight_ih_l0"
K_WHH = "sdf_net.macro_lstm.lstm.weight_hh_l0"
K_BIH = "sdf_net.macro_lstm.lstm.bias_ih_l0"
K_BHH = "sdf_net.macro_lstm.lstm.bias_hh_l0"
K_W1 = "sdf_net.fc_layers.0.weight"
K_B1 = "sdf_net.fc_layers.0.bias"
K_W2 = "sdf_net.fc_layers.3.weight"
K_B2 = "sdf_net.fc_layers.3.bias"
K_W3 = "sdf_net.output_proj.weight"
K_B3 = "sdf_net.output_proj.bias"
K_WQ = "moment_net.output_proj.weight"
K_BQ = "moment_net.output_proj.bias"
PARAM_KEYS = (K_WIH, K_WHH, K_BIH, K_BHH, K_W1, K_B1, K_W2, K_B2, K_W3, K_B3, K_WQ, K_BQ)
SDF_KEYS = PARAM_KEYS[:10]
MOMENT_KEYS = PARAM_KEYS[10:]


def init_params(macro_dim: int, char_dim: int, seed: int, rnn_units: int = 4,
                hidden=(64, 64), num_moments: int = 8,
                dtype: torch.dtype = torch.float32) -> Dict[str, torch.Tensor]:
    """Default-initialised parameters, identical to ``torch.manual_seed(seed);
    AssetPricingGAN(cfg)`` because the torch modules are created in the reference's order:
    LSTM (src/model.py:39), the hidden Linears (:211), the SDF output Linear (:219), then the
    moment Linear (:130).  ReLU/Dropout consume no random numbers."""
    torch.manual_seed(seed)
    lstm = torch.nn.LSTM(macro_dim, rnn_units, 1, batch_first=True)
    fc0 = torch.nn.Linear(char_dim + rnn_units, hidden[0])
    fc3 = torch.nn.Linear(hidden[0], hidden[1])
    out = torch.nn.Linear(hidden[1], 1)
    mom = torch.nn.Linear(macro_dim + char_dim, num_moments)
    p = {
        K_WIH: lstm.weight_ih_l0, K_WHH: lstm.weight_hh_l0,
        K_BIH: lstm.bias_ih_l0, K_BHH: lstm.bias_hh_l0,
        K_W1: fc0.weight, K_B1: fc0.bias, K_W2: fc3.weight, K_B2: fc3.bias,
        K_W3: out.weight, K_B3: out.bias, K_WQ: mom.weight, K_BQ: mom.bias,
    }
    return {k: v.detach().clone().to(dtype) for k, v in p.items()}


def macro_lstm(params, macro: torch.Tensor, hidden=None):
    """MacroLSTM.forward (src/model.py:50-73): one-layer LSTM over a batch of one sequence,
    gate order i,f,g,o, zero initial state when ``hidden`` is None."""
    seq = macro.unsqueeze(0)
    hsz = params[K_WHH].shape[1]
    if hidden is None:
        z = torch.zeros(1, 1, hsz, dtype=macro.dtype, device=macro.device)
        hidden = (z, z.clone())
    flat = [params[K_WIH], params[K_WHH], params[K_BIH], params[K_BHH]]
    out, hn, cn = torch._VF.lstm(seq, hidden, flat, True, 1, 0.0, False, False, True)
    return out.squeeze(0), (hn, cn)


def sdf_weights(params, macro, x, mask, hidden=None, normalize_w: bool = True):
    """SDFNetwork.forward (src/model.py:221-281)."""
    T, N, _ = x.shape
    new_hidden = None
    if macro is not None:
        state, new_hidden = macro_lstm(params, macro, hidden)
        cells = torch.cat([x, state.unsqueeze(1).expand(-1, N, -1)], dim=-1)  # :253-255
    else:
        cells = x
    flat = cells.view(T * N, -1)
    h = F.relu(F.linear(flat, params[K_W1], params[K_B1]))                     # :211-212
    h = F.relu(F.linear(h, params[K_W2], params[K_B2]))
    w = F.linear(h, params[K_W3], params[K_B3]).view(T, N)                     # :267-268
    m = mask.float()
    w = w * m                                                                  # :271
    if normalize_w:                                                            # :274-279
        tot = (w * m).sum(dim=1, keepdim=True)
        cnt = m.sum(dim=1, keepdim=True).clamp(min=1)
        w = (w - tot / cnt) * m
    return w, new_hidden


def moment_conditions(params, macro, x):
    """Moment-net input build (src/model.py:514-518) and MomentNetwork.forward (:132-161) for
    the default ``hidden_dim_moment=[]``: tanh(Linear([macro, x])) as a [K,T,N] permuted view."""
    T, N, _ = x.shape
    if macro is not None:
        inp = torch.cat([macro.unsqueeze(1).expand(-1, N, -1), x], dim=-1)
    else:
        inp = x
    q = torch.tanh(F.linear(inp.view(T * N, -1), params[K_WQ], params[K_BQ]))
    return q.view(T, N, -1).permute(2, 0, 1)


def _portfolio(w, R, mask, weighted_loss):
    """Shared head of both loss functions (src/model.py:359-369 and :403-413)."""
    m = mask.float()
    n_t = m.sum(dim=1, keepdim=True).clamp(min=1)
    n_bar = n_t.mean()
    wr = w * R * m
    if weighted_loss:
        pr = wr.sum(dim=1) / n_t.squeeze() * n_bar
    else:
        pr = wr.sum(dim=1)
    return pr


def unconditional_loss(w, R, mask, weighted_loss: bool = True):
    """compute_unconditional_loss (src/model.py:346-387)."""
    pr = _portfolio(w, R, mask, weighted_loss)
    m = mask.float()
    t_i = m.sum(dim=0).clamp(min=1)
    sdf = pr.unsqueeze(1) + 1                                                  # :380
    e = (R * m * sdf).sum(dim=0) / t_i                                         # :382
    return (e ** 2).mean(), pr


def conditional_loss(w, R, mask, q, weighted_loss: bool = True):
    """compute_conditional_loss (src/model.py:389-433), including its loop over moments."""
    pr = _portfolio(w, R, mask, weighted_loss)
    m = mask.float()
    sdf = pr.unsqueeze(1) + 1
    t_i = m.sum(dim=0).clamp(min=1)
    per_k = []
    for k in range(q.shape[0]):                                                # :425-429
        e = (q[k] * R * m * sdf).sum(dim=0) / t_i
        per_k.append((e ** 2).mean())
    return torch.stack(per_k).mean(), pr


def residual_loss(w, R, mask):
    """compute_residual_loss (src/model.py:435-483): per month with at least two listed stocks, the mean squared
    residual of the returns after projecting them on the weights (months whose weights vanish are left out of the
    numerator only), over the mean squared return."""
    res, rsq = [], []
    for t in range(w.shape[0]):
        valid = mask[t].float() > 0                                            # :455-459
        if valid.sum() < 2:
            continue
        wv, rv = w[t][valid], R[t][valid]
        ww = (wv * wv).sum()                                                   # :464-470
        if ww > 1e-8:
            rhat = ((rv * wv).sum() / ww) * wv
            res.append(((rv - rhat) ** 2).mean())
        rsq.append((rv ** 2).mean())                                           # :472
    if len(res) == 0:                                                          # :474-475
        return torch.tensor(0.0, device=w.device)
    return torch.stack(res).mean() / torch.stack(rsq).mean().clamp(min=1e-8)   # :477-480


def forward(params, macro, x, R, mask, hidden=None, phase: str = "conditional",
            weighted_loss: bool = True, normalize_w: bool = True,
            residual_loss_factor: float = 0.0) -> Dict[str, torch.Tensor]:
    """AssetPricingGAN.forward (src/model.py:485-563)."""
    w, new_hidden = sdf_weights(params, macro, x, mask, hidden, normalize_w)
    q = moment_conditions(params, macro, x)
    zero = torch.zeros((), device=x.device)
    if phase == "unconditional":                                               # :524-528
        l_u, pr = unconditional_loss(w, R, mask, weighted_loss)
        l_c, total = zero, l_u
    elif phase == "moment":                                                    # :530-535
        l_c, pr = conditional_loss(w, R, mask, q, weighted_loss)
        l_u, total = zero, -l_c
    else:                                                                      # :537-541
        l_c, pr = conditional_loss(w, R, mask, q, weighted_loss)
        l_u, _ = unconditional_loss(w, R, mask, weighted_loss)
        total = l_c
    l_res = zero
    if residual_loss_factor > 0:                                               # :543-548
        l_res = residual_loss(w, R, mask)
        total = total + residual_loss_factor * l_res
    sharpe = pr.mean() / (pr.std() + 1e-8)                                     # :551
    return {"weights": w, "loss": total, "loss_unconditional": l_u, "loss_conditional": l_c,
            "loss_residual": l_res, "sharpe": sharpe, "portfolio_returns": pr,
            "hidden": new_hidden, "moments": q}


def normalized_weights(params, macro, x, mask, hidden=None):
    """get_weights(normalized=True) (src/model.py:565-594): each month divided by
    clamp(sum |w|*m, 1e-8), month by month as the reference does."""
    w, new_hidden = sdf_weights(params, macro, x, mask, hidden)
    out = torch.zeros_like(w)
    for t in range(w.shape[0]):
        s = (w[t].abs() * mask[t].float()).sum().clamp(min=1e-8)
        out[t] = w[t] / s
    return out, new_hidden


def step(params, macro, x, R, mask, phase: str = "conditional", hidden=None,
         grad_keys: Optional[Tuple[str, ...]] = None, **forward_kw):
    """One forward + backward the way train_epoch drives it (src/train.py:77-84): returns the
    forward dict (detached) and the gradient of ``loss`` for every key in ``grad_keys``
    (default: all twelve; the caller freezes groups by passing a subset, mirroring
    ``requires_grad=False`` at src/train.py:309-312)."""
    keys = PARAM_KEYS if grad_keys is None else tuple(grad_keys)
    live = {k: (v.detach().clone().requires_grad_(k in keys)) for k, v in params.items()}
    out = forward(live, macro, x, R, mask, hidden=hidden, phase=phase, **forward_kw)
    out["loss"].backward()
    grads = {k: (live[k].grad if live[k].grad is not None else None) for k in keys}
    det = {}
    for k, v in out.items():
        if isinstance(v, torch.Tensor):
            det[k] = v.detach()
        elif isinstance(v, tuple):
            det[k] = tuple(a.detach() for a in v)
        else:
            det[k] = v
    return det, grads
