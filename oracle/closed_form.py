"""numpy float64 closed-form restatement (forward AND backward) of the adversarial-SDF step.
TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py`` for who may import this).

Parity status: PINNED -- ``tests/test_oracle_golden.py`` checks the forward values and all
twelve gradients against fixtures written by the reference itself
(``tests/golden/make_golden.py``), at float32-rounding tolerance (the fixtures are the
reference's fp32 results; this file works in fp64).

Unlike ``torch_port`` this file does not use autograd: the backward pass is the hand-derived
chain of SURVEY.md §8 a9, which is exactly what the CUDA kernels implement, stage by stage and
under the same names.  Citations are ``/root/reference`` paths.

Conventions (defaults of src/model.py:304-344): F characteristics, M macros, H LSTM units,
D1/D2 hidden widths, K moments.  ``W1 = [W1x | W1m]`` (characteristics first, src/model.py:255),
``Wq = [Wqm | Wqx]`` (macro first, src/model.py:516).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np

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


def _sig(v):
    return 1.0 / (1.0 + np.exp(-v))


def lstm_forward(p, macro, h0=None, c0=None):
    """torch.nn.LSTM semantics used at src/model.py:39-45,68: gates i,f,g,o;
    c_t = f*c + i*g ; h_t = o*tanh(c_t).  Returns per-step h and everything BPTT needs."""
    T = macro.shape[0]
    H = p[K_WHH].shape[1]
    h = np.zeros(H) if h0 is None else np.asarray(h0, dtype=np.float64).reshape(H)
    c = np.zeros(H) if c0 is None else np.asarray(c0, dtype=np.float64).reshape(H)
    xproj = macro @ p[K_WIH].T + p[K_BIH] + p[K_BHH]
    hs = np.zeros((T, H)); cs = np.zeros((T, H)); gates = np.zeros((T, 4 * H))
    hprev = np.zeros((T, H)); cprev = np.zeros((T, H))
    for t in range(T):
        hprev[t], cprev[t] = h, c
        pre = xproj[t] + p[K_WHH] @ h
        i, f, g, o = _sig(pre[:H]), _sig(pre[H:2 * H]), np.tanh(pre[2 * H:3 * H]), _sig(pre[3 * H:])
        c = f * c + i * g
        h = o * np.tanh(c)
        hs[t], cs[t] = h, c
        gates[t] = np.concatenate([i, f, g, o])
    return hs, dict(gates=gates, cs=cs, hprev=hprev, cprev=cprev, hn=h, cn=c)


def lstm_backward(p, macro, saved, dh_ext):
    """BPTT for the recurrence above, given dL/dh_t injected at every step."""
    T, H = dh_ext.shape
    gates, cs, hprev, cprev = saved["gates"], saved["cs"], saved["hprev"], saved["cprev"]
    dpre = np.zeros((T, 4 * H))
    dh_rec = np.zeros(H); dc = np.zeros(H)
    for t in range(T - 1, -1, -1):
        i, f, g, o = (gates[t, :H], gates[t, H:2 * H], gates[t, 2 * H:3 * H], gates[t, 3 * H:])
        dh = dh_ext[t] + dh_rec
        tc = np.tanh(cs[t])
        dc = dc + dh * o * (1 - tc * tc)
        dpre[t, :H] = dc * g * i * (1 - i)
        dpre[t, H:2 * H] = dc * cprev[t] * f * (1 - f)
        dpre[t, 2 * H:3 * H] = dc * i * (1 - g * g)
        dpre[t, 3 * H:] = dh * tc * o * (1 - o)
        dh_rec = p[K_WHH].T @ dpre[t]
        dc = dc * f
    db = dpre.sum(axis=0)
    return {K_WIH: dpre.T @ macro, K_WHH: dpre.T @ hprev, K_BIH: db, K_BHH: db.copy()}


def forward(params: Dict[str, np.ndarray], macro, x, R, mask, h0=None, c0=None,
            phase: str = "conditional", weighted_loss: bool = True) -> Dict[str, np.ndarray]:
    """Forward of AssetPricingGAN.forward (src/model.py:485-563), restructured the way the
    kernels compute it: the two torch.cat's (src/model.py:255, :516) become per-month biases
    ``a_t`` / ``c_t``; demeaning (:274-279) and the portfolio return (:363-367) come from the
    four row sums; the loss (:376-385, :418-431) from per-asset column sums."""
    p = {k: np.asarray(v, dtype=np.float64) for k, v in params.items()}
    x = np.asarray(x, dtype=np.float64); R = np.asarray(R, dtype=np.float64)
    m = np.asarray(mask).astype(np.float64)
    T, N, F = x.shape
    K = p[K_WQ].shape[0]
    s: Dict[str, np.ndarray] = {}
    if macro is not None:
        macro = np.asarray(macro, dtype=np.float64)
        hs, saved = lstm_forward(p, macro, h0, c0)
        a = hs @ p[K_W1][:, F:].T + p[K_B1]                       # [T,D1]
        c = macro @ p[K_WQ][:, :macro.shape[1]].T + p[K_BQ]       # [T,K]
        wqx = p[K_WQ][:, macro.shape[1]:]
        s.update(hs=hs, lstm=saved)
    else:
        a = np.broadcast_to(p[K_B1], (T, p[K_B1].shape[0]))
        c = np.broadcast_to(p[K_BQ], (T, K))
        wqx = p[K_WQ]
    z1 = x @ p[K_W1][:, :F].T + a[:, None, :]
    h1 = np.maximum(z1, 0)
    z2 = h1 @ p[K_W2].T + p[K_B2]
    h2 = np.maximum(z2, 0)
    u = h2 @ p[K_W3][0] + p[K_B3][0]                              # [T,N]
    q = np.tanh(x @ wqx.T + c[:, None, :])                        # [T,N,K]
    n_t = np.maximum(m.sum(axis=1), 1.0)                          # src/model.py:277, :359
    mu = (m * u).sum(axis=1) / n_t
    w = m * (u - mu[:, None])                                     # == :271-279
    n_bar = n_t.mean()
    P = (w * R * m).sum(axis=1)
    pr = P / n_t * n_bar if weighted_loss else P
    S = 1.0 + pr                                                  # :380
    t_i = np.maximum(m.sum(axis=0), 1.0)                          # :376
    rm = R * m
    e0 = (rm * S[:, None]).sum(axis=0) / t_i                      # :382
    e = np.einsum("tnk,tn->kn", q, rm * S[:, None]) / t_i         # :428
    l_u = (e0 ** 2).mean()
    l_c = (e ** 2).mean(axis=1).mean()                            # :429-431
    if phase == "unconditional":
        loss, l_c_rep, l_u_rep = l_u, 0.0, l_u
    elif phase == "moment":
        loss, l_c_rep, l_u_rep = -l_c, l_c, 0.0
    else:
        loss, l_c_rep, l_u_rep = l_c, l_c, l_u
    sharpe = pr.mean() / (pr.std(ddof=1) + 1e-8) if T > 1 else np.nan   # :551 (unbiased)
    s.update(a=a, c=c, z1=z1, h1=h1, z2=z2, h2=h2, u=u, q=q, n_t=n_t, mu=mu, n_bar=n_bar,
             S=S, t_i=t_i, e0=e0, e=e, m=m, rm=rm, wqx=wqx)
    return {"weights": w, "loss": loss, "loss_unconditional": l_u_rep,
            "loss_conditional": l_c_rep, "sharpe": sharpe, "portfolio_returns": pr,
            "moments": np.transpose(q, (2, 0, 1)), "_saved": s,
            "hidden": (s["lstm"]["hn"], s["lstm"]["cn"]) if macro is not None else None}


def backward(params, macro, x, R, mask, out, phase: str = "conditional",
             weighted_loss: bool = True) -> Dict[str, np.ndarray]:
    """d loss / d params for all twelve tensors.  ``out`` is what
    ``forward`` returned.  Gradients of groups the reference would have frozen are simply
    ignored by the caller."""
    p = {k: np.asarray(v, dtype=np.float64) for k, v in params.items()}
    x = np.asarray(x, dtype=np.float64)
    s = out["_saved"]
    T, N, F = x.shape
    K = p[K_WQ].shape[0]
    m, rm, S, t_i, q = s["m"], s["rm"], s["S"], s["t_i"], s["q"]
    if phase == "unconditional":
        G = np.zeros((K, N)); G0 = 2.0 * s["e0"] / (N * t_i)
    else:
        sign = -1.0 if phase == "moment" else 1.0
        G = sign * 2.0 * s["e"] / (K * N * t_i); G0 = np.zeros(N)
    # dL/dS_t, then through pr_t = P_t * n_bar / n_t and P_t = sum_i m (u - mu_t) R
    gq = np.einsum("kn,tnk->tn", G, q) + G0[None, :]
    dS = (rm * gq).sum(axis=1)
    dP = dS * (s["n_bar"] / s["n_t"]) if weighted_loss else dS
    rbar = rm.sum(axis=1) / s["n_t"]
    du = m * dP[:, None] * (np.asarray(R, dtype=np.float64) - rbar[:, None])
    # SDF MLP
    g = {}
    g[K_W3] = np.einsum("tn,tnj->j", du, s["h2"])[None, :]
    g[K_B3] = np.array([du.sum()])
    dz2 = du[:, :, None] * p[K_W3][0][None, None, :] * (s["z2"] > 0)
    g[K_W2] = np.einsum("tnj,tnk->jk", dz2, s["h1"])
    g[K_B2] = dz2.sum(axis=(0, 1))
    dz1 = (dz2 @ p[K_W2]) * (s["z1"] > 0)
    dw1x = np.einsum("tnj,tnf->jf", dz1, x)
    da = dz1.sum(axis=1)                                           # [T,D1]
    # moment net
    dq = G.T[None, :, :] * (rm * S[:, None])[:, :, None]           # [T,N,K]
    dpre = dq * (1 - q * q)
    dwqx = np.einsum("tnk,tnf->kf", dpre, x)
    dc = dpre.sum(axis=1)                                          # [T,K]
    if macro is not None:
        macro = np.asarray(macro, dtype=np.float64)
        g[K_W1] = np.concatenate([dw1x, da.T @ s["hs"]], axis=1)
        g[K_B1] = da.sum(axis=0)
        dh_ext = da @ p[K_W1][:, F:]
        g.update(lstm_backward(p, macro, s["lstm"], dh_ext))
        g[K_WQ] = np.concatenate([dc.T @ macro, dwqx], axis=1)
    else:
        g[K_W1] = dw1x
        g[K_B1] = da.sum(axis=0)
        g[K_WQ] = dwqx
    g[K_BQ] = dc.sum(axis=0)
    return g


def normalized_weights(w, mask):
    """get_weights(normalized=True) tail (src/model.py:581-592)."""
    m = np.asarray(mask).astype(np.float64)
    s = np.maximum((np.abs(w) * m).sum(axis=1), 1e-8)
    return w / s[:, None]


def mask_from_raw(ret, feat, missing: float = -99.99):
    """Mask and zero-fill contract of the loader (src/data_loader.py:50-65), fp32 inputs.
    Bit-exact: integer/boolean logic only."""
    ret = np.asarray(ret, dtype=np.float32); feat = np.asarray(feat, dtype=np.float32)
    thr = missing + 1
    mask = (ret > thr) & ~np.isnan(ret) & np.all(feat > thr, axis=2)
    return mask, np.where(mask, ret, 0.0).astype(np.float32), \
        np.where(mask[:, :, None], feat, 0.0).astype(np.float32)


def _numpy_pairwise(col):
    """numpy's pairwise summation of one contiguous fp32 run (the scalar loop behind add.reduce along the contiguous
    axis): fewer than 8 terms one after the other; up to 128 terms eight interleaved accumulators, combined as
    ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), then the remainder; longer runs split at n/2 rounded down to a multiple of 8."""
    f = np.float32
    n = len(col)
    if n < 8:
        r = f(0)
        for v in col:
            r = f(r + v)
        return r
    if n <= 128:
        r = [f(col[j]) for j in range(8)]
        i = 8
        while i < n - (n % 8):
            for j in range(8):
                r[j] = f(r[j] + col[i + j])
            i += 8
        res = f(f(f(r[0] + r[1]) + f(r[2] + r[3])) + f(f(r[4] + r[5]) + f(r[6] + r[7])))
        while i < n:
            res = f(res + col[i]); i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return f(_numpy_pairwise(col[:n2]) + _numpy_pairwise(col[n2:]))


def macro_zscore(macro, mean=None, std=None, pairwise=False):
    """Macro normalisation of the loader (src/data_loader.py:76-86) with the summation order written out.
    pairwise=False: the file's array as loaded, C-contiguous [T, M]; numpy reduces it along axis 0 by adding the rows
    one after the other in fp32.  pairwise=True: after ``macro_features[:, macro_idx]`` (src/data_loader.py:72-73)
    the array is COLUMN-major (fancy indexing on the last axis), axis 0 is the contiguous one, and numpy sums each
    column with its pairwise scheme.  `std` subtracts the mean, squares, sums the same way, divides by T and takes the
    root; the loader adds 1e-8.  Returns (normalized, mean [1,M], std [1,M]), all fp32, bit-identical to the
    reference's arrays (tests/test_oracle_golden.py) -- which is what the device kernel is held to."""
    a = np.ascontiguousarray(macro, dtype=np.float32)
    T, M = a.shape

    def colsum(v):
        if pairwise:
            return np.array([_numpy_pairwise(v[:, m]) for m in range(M)], np.float32)
        s = np.zeros(M, np.float32)
        for t in range(T):
            s = s + v[t]
        return s

    if mean is None:
        mean = (colsum(a) / np.float32(T))[None]
        d = a - mean
        std = (np.sqrt(colsum(d * d) / np.float32(T)) + np.float32(1e-8))[None]
    mean = np.asarray(mean, np.float32).reshape(1, M); std = np.asarray(std, np.float32).reshape(1, M)
    return (a - mean) / std, mean, std


def small_sample(mask, n_periods, n_stocks):
    """Row/column selection of create_small_sample (src/data_loader.py:212-237): months [0, T'), the N' stocks with
    the most valid months in numpy's argsort order.  Returns (T', stock indices)."""
    T = min(n_periods, mask.shape[0]); N = min(n_stocks, mask.shape[1])
    return T, np.argsort(mask.sum(axis=0))[-N:]

