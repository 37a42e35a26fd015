"""GPU parity tests proper: the CUDA path (through the C ABI) against the oracle and the goldens
written by the reference itself.  Tolerances are BASELINE.json's: 1e-5 relative on losses,
1e-4 on weights (relative to the largest weight), mask/index logic exact."""
import ctypes as C

import numpy as np
import pytest
import torch

from _dlap_testlib import PHASES, golden_params, load_golden, rel_err
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, _lib
from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
from oracle import closed_form as cf

pytestmark = pytest.mark.gpu
LOSS_RTOL = 1e-5
WEIGHT_TOL = 1e-4
GRAD_TOL = 2e-4     # fp32 path: relative to the largest entry of each gradient tensor


def dev():
    return torch.device("cuda:0")


def _model_from(params, M, F=46, **cfg):
    cfg.setdefault("b200_precision", "fp32")      # this file pins the exact-fp32 path unless a test says otherwise
    model = AssetPricingGAN({"macro_feature_dim": M, "individual_feature_dim": F, **cfg})
    model.load_state_dict({k: torch.as_tensor(v) for k, v in params.items()})
    return model.to(dev())


def _panel(g):
    d = dev()
    hidden = None
    if "h0" in g:
        hidden = (torch.from_numpy(g["h0"]).to(d), torch.from_numpy(g["c0"]).to(d))
    return (torch.from_numpy(g["macro"]).to(d), torch.from_numpy(g["x"]).to(d),
            torch.from_numpy(g["R"]).to(d), torch.from_numpy(g["mask"]).to(d), hidden)


def make_panel(T, N, M=178, F=46, seed=0, valid=0.4):
    g = torch.Generator().manual_seed(seed)
    mask = torch.rand(T, N, generator=g) < valid
    x = torch.randn(T, N, F, generator=g).clamp_(-2, 2) * mask[:, :, None]
    R = 0.1 * torch.randn(T, N, generator=g) * mask
    macro = torch.randn(T, M, generator=g) if M else None
    return macro, x, R, mask


def _np(t):
    return None if t is None else t.detach().cpu().numpy()


def _check_against_cf(model, macro, x, R, mask, phase, hidden=None, grad_tol=GRAD_TOL):
    d = dev()
    params = {k: _np(v) for k, v in model.state_dict().items()}
    model.zero_grad(set_to_none=True)
    frozen = phase == "moment"
    for p in model.sdf_net.parameters():
        p.requires_grad = not frozen
    to = lambda t: None if t is None else t.to(d)
    hid = None if hidden is None else (hidden[0].to(d), hidden[1].to(d))
    out = model(to(macro), to(x), to(R), to(mask), hidden=hid, phase=phase)
    out["loss"].backward()
    torch.cuda.synchronize()
    h0 = None if hidden is None else _np(hidden[0]).reshape(-1)
    c0 = None if hidden is None else _np(hidden[1]).reshape(-1)
    ref = cf.forward(params, _np(macro), _np(x), _np(R), _np(mask), h0=h0, c0=c0, phase=phase)
    gref = cf.backward(params, _np(macro), _np(x), _np(R), _np(mask), ref, phase=phase)
    w = _np(out["weights"])
    assert np.all(w[~_np(mask)] == 0.0), "masked cells must be exactly zero"
    assert rel_err(w, ref["weights"]) < WEIGHT_TOL
    assert rel_err(_np(out["portfolio_returns"]), ref["portfolio_returns"]) < WEIGHT_TOL
    for key in ("loss", "loss_unconditional", "loss_conditional"):
        a, b = float(out[key]), float(ref[key])
        assert abs(a - b) <= LOSS_RTOL * abs(b) + 1e-30, (key, a, b)
    if x.shape[0] > 1:
        assert abs(float(out["sharpe"]) - float(ref["sharpe"])) <= 1e-4 * abs(float(ref["sharpe"])) + 1e-7
    for k, p in model.named_parameters():
        is_mom = k.startswith("moment_net")
        expect = (is_mom and phase != "unconditional") or (not is_mom and phase != "moment")
        if not expect:
            assert p.grad is None, k
            continue
        assert p.grad is not None, k
        r = gref[k].reshape(p.shape)
        if k == "sdf_net.output_proj.bias":
            assert float(p.grad.abs().max()) < 1e-9     # identically zero (demeaning)
            continue
        scale = max(np.abs(r).max(), 1e-30)
        err = np.abs(_np(p.grad) - r).max() / scale
        assert err < grad_tol, (k, err)
    for p in model.sdf_net.parameters():
        p.requires_grad = True
    return out, ref


# ---- stage-level checks through the raw C ABI ------------------------------------------------------
def test_macro_fwd_stage():
    g = load_golden("rand_small")
    params = golden_params(g)
    model = _model_from(params, 178)
    macro = torch.from_numpy(g["macro"]).to(dev())
    out, (hn, cn) = model.sdf_net.macro_lstm(macro)
    hs, saved = cf.lstm_forward({k: v.astype(np.float64) for k, v in params.items()}, g["macro"].astype(np.float64))
    assert np.abs(_np(out) - hs).max() < 2e-6
    assert np.abs(_np(hn).reshape(-1) - saved["hn"]).max() < 2e-6
    assert np.abs(_np(cn).reshape(-1) - saved["cn"]).max() < 2e-6


def _check_reduce_stage(red, ref, s, T, N):
    """dlap_step_reduce against the closed form: per-month statistics, losses, unit loss coefficients, dS sums."""
    rs = _np(red.row_stats)
    m = s["m"]
    assert np.array_equal(rs[:, Fn.RS_NT], np.maximum(m.sum(axis=1), 1))           # counts are exact integers
    assert np.abs(rs[:, Fn.RS_MU] - s["mu"]).max() < 1e-6 * max(1.0, np.abs(s["mu"]).max())
    assert np.abs(rs[:, Fn.RS_S] - s["S"]).max() < 1e-6
    assert np.abs(_np(red.pr) - ref["portfolio_returns"]).max() < 1e-4 * np.abs(ref["portfolio_returns"]).max()
    sc = _np(red.scalars)
    assert abs(sc[Fn.SC_LC] - ref["loss_conditional"]) <= LOSS_RTOL * ref["loss_conditional"]
    assert abs(sc[Fn.SC_LU] - ref["loss_unconditional"]) <= LOSS_RTOL * ref["loss_unconditional"]
    assert abs(sc[Fn.SC_NBAR] - s["n_bar"]) <= 1e-6 * s["n_bar"]
    assert abs(sc[Fn.SC_SHARPE] - ref["sharpe"]) <= 1e-4 * abs(ref["sharpe"]) + 1e-7
    assert rel_err(_np(red.w), ref["weights"]) < WEIGHT_TOL
    K = 8
    Gc = 2.0 * s["e"] / (K * N * s["t_i"]); Gu = 2.0 * s["e0"] / (N * s["t_i"])
    G = _np(red.G).T                                           # [N,12] stock-major rows -> [12,N]
    assert np.abs(G[:8] - Gc).max() < 1e-5 * np.abs(Gc).max() and np.abs(G[8] - Gu).max() < 1e-5 * np.abs(Gu).max()
    assert not G[9:].any()
    dSc = (s["rm"] * np.einsum("kn,tnk->tn", Gc, s["q"])).sum(axis=1)
    dSu = (s["rm"] * Gu[None, :]).sum(axis=1)
    dS2 = _np(red.dS2)
    assert np.abs(dS2[:, 0] - dSc).max() < 2e-5 * np.abs(dSc).max() and np.abs(dS2[:, 1] - dSu).max() < 2e-5 * np.abs(dSu).max()


def test_forward_stages_rand_small():
    g = load_golden("rand_small")
    params = golden_params(g)
    model = _model_from(params, 178).eval()
    macro, x, R, mask, _ = _panel(g)
    T, N, F = x.shape
    ref = cf.forward(params, g["macro"], g["x"], g["R"], g["mask"])
    s = ref["_saved"]
    kp = [p.detach() for p in model._inputs(macro, x, True).params]
    ws = Fn.workspace(T, N, F, 178, 4, dev())
    cp = Fn.cell_forward(ws, macro, x, mask, kp, None, None, "fp32", True)
    torch.cuda.synchronize()
    assert np.abs(_np(cp.a_t) - s["a"]).max() < 5e-6
    assert np.abs(_np(cp.c_t) - s["c"]).max() < 2e-5
    assert np.abs(_np(cp.u) - s["u"]).max() < 1e-5 * max(1.0, np.abs(s["u"]).max())
    assert np.abs(_np(cp.q) - np.transpose(s["q"], (2, 0, 1))).max() < 1e-6          # dense, masked cells included
    red = Fn.step_reduce(ws, cp, R, mask, True, True)
    torch.cuda.synchronize()
    _check_reduce_stage(red, ref, s, T, N)
    assert int(ws.bar.abs().sum()) == 0                       # the barrier words are back to zero


def _index_layout(T, N):
    off_ts, off_cnt = 4, 4 + T + 1
    off_vidx = (4 + 2 * T + 1 + 3) & ~3
    return off_ts, off_cnt, off_vidx, off_vidx + T * N


@pytest.mark.parametrize("shape,density", [((7, 1000), 0.4), ((5, 130), 0.0), ((3, 257), 1.0), ((240, 3000), 0.35),
                                           ((2, 1), 0.5), ((1100, 40), 0.5)])
def test_panel_index_bit_exact(shape, density):
    """dlap_panel_index against numpy: counts, tile prefix and the ascending stock lists, bit for bit."""
    T, N = shape
    rng = np.random.default_rng(T * 1000 + N)
    mask = rng.random((T, N)) < density
    if T > 2 and 0.0 < density < 1.0:
        mask[1] = False                              # an empty month in the middle
    ws = Fn.workspace(T, N, 46, 0, 0, dev())
    idx = _np(ws.panel_index(torch.from_numpy(mask).to(dev())))
    off_ts, off_cnt, off_vidx, off_rank = _index_layout(T, N)
    cnt = mask.sum(axis=1)
    tiles = (cnt + 127) // 128
    ts = np.concatenate([[0], np.cumsum(tiles)])
    assert idx[0] == ts[-1]
    assert np.array_equal(idx[off_ts:off_ts + T + 1], ts)
    assert np.array_equal(idx[off_cnt:off_cnt + T], cnt)
    vidx = idx[off_vidx:off_vidx + T * N].reshape(T, N)
    for t in range(T):
        assert np.array_equal(vidx[t, :cnt[t]], np.nonzero(mask[t])[0])
    nw = (N + 31) // 32
    rank = idx[off_rank:off_rank + T * nw].reshape(T, nw)
    before = np.concatenate([np.zeros((T, 1), dtype=np.int64), np.cumsum(mask, axis=1)], axis=1)[:, 0:N:32]
    assert np.array_equal(rank, before)
    off_tm = (off_rank + T * nw + 3) & ~3                # month of every tile
    assert np.array_equal(idx[off_tm:off_tm + ts[-1]], np.repeat(np.arange(T), tiles))
    assert idx.size >= off_tm + T * ((N + 127) // 128)


def test_forward_stages_tensor_core_visits_only_unmasked_cells():
    """BF16X3 forward through the ABI: u and q are written for unmasked cells only, left-packed per month
    (poisoned buffers keep their poison elsewhere); the LSTM that runs on one CTA of the same grid leaves the same
    tables as the stand-alone macro kernel; the reductions on the packed layout match the closed form."""
    g = load_golden("rand_small")
    params = golden_params(g)
    model = _model_from(params, 178).eval()
    macro, x, R, mask, _ = _panel(g)
    T, N, F = x.shape
    ref = cf.forward(params, g["macro"], g["x"], g["R"], g["mask"])
    s = ref["_saved"]
    kp = [p.detach() for p in model._inputs(macro, x, True).params]
    ws = Fn.workspace(T, N, F, 178, 4, dev())
    save, a_t, c_t, hn, cn = Fn.macro_forward(ws, macro, kp, None, None)            # stand-alone macro kernel
    real_empty = torch.empty
    try:                                                      # poison what cell_forward allocates
        torch.empty = lambda *a, **k: real_empty(*a, **k).fill_(float("nan")) if k.get("dtype") == torch.float32 else real_empty(*a, **k)
        cp = Fn.cell_forward(ws, macro, x, mask, kp, None, None, "bf16x3", True)
    finally:
        torch.empty = real_empty
    torch.cuda.synchronize()
    assert torch.equal(cp.a_t, a_t) and torch.equal(cp.c_t, c_t) and torch.equal(cp.save, save)
    assert torch.equal(cp.hn, hn) and torch.equal(cp.cn, cn)
    assert [int(v) for v in ws.sync] == [0, 0]                # progress words are back to zero
    m = g["mask"]
    un, qn = _np(cp.u), _np(cp.q)
    qref = np.transpose(s["q"], (2, 0, 1))
    for t in range(T):                                   # the j-th unmasked stock of month t sits at [t, j]
        c = int(m[t].sum())
        assert np.isnan(un[t, c:]).all() and np.isnan(qn[:, t, c:]).all()
        assert np.abs(un[t, :c] - s["u"][t, m[t]]).max(initial=0.0) < 1e-5 * max(1.0, np.abs(s["u"]).max())
        assert np.abs(qn[:, t, :c] - qref[:, t, m[t]]).max(initial=0.0) < 1e-5
    red = Fn.step_reduce(ws, cp, R, mask, True, True)
    torch.cuda.synchronize()
    _check_reduce_stage(red, ref, s, T, N)


@pytest.mark.parametrize("shape", [(7, 1000, 46), (3, 130, 45), (2, 257, 46)])
def test_upload_panel_pulls_exactly_the_unmasked_rows(shape):
    """upload_panel from pinned host memory (dlap_panel_pull): unmasked rows bit-identical to the host panel,
    masked rows zero; returns, mask and macro series copied whole.  A step on the pulled panel is bit-identical
    to a step on a plain .to(device) copy."""
    from deeplearninginassetpricing_paperreplication_b200 import upload_panel
    T, N, F = shape
    rng = np.random.default_rng(5)
    mask = torch.from_numpy(rng.random((T, N)) < 0.4)
    mask[1] = False
    x = torch.from_numpy(rng.standard_normal((T, N, F)).astype(np.float32))       # NOT zero-filled where masked
    R = torch.from_numpy((0.1 * rng.standard_normal((T, N))).astype(np.float32)) * mask
    macro = torch.from_numpy(rng.standard_normal((T, 178)).astype(np.float32))
    host = [t.pin_memory() for t in (macro, x, R, mask)]
    d_macro, d_x, d_R, d_mask = upload_panel(*host, device=dev())
    torch.cuda.synchronize()
    assert torch.equal(d_mask.cpu(), mask) and torch.equal(d_R.cpu(), R) and torch.equal(d_macro.cpu(), macro)
    xd = d_x.cpu()
    assert torch.equal(xd[mask], x[mask])
    assert float(xd[~mask].abs().max()) == 0.0
    if F == 46:
        outs = []
        for xin in (d_x, (x * mask[:, :, None]).to(dev())):
            torch.manual_seed(3)
            model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": F, "dropout": 0.0,
                                     "b200_precision": "bf16x3"}).to(dev())
            out = model(d_macro, xin, d_R, d_mask)
            out["loss"].backward()
            outs.append((out["loss"].detach().clone(), out["weights"].detach().clone(),
                         model.sdf_net.fc_layers[0].weight.grad.clone()))
        for a, b in zip(*outs):
            assert torch.equal(a, b)


def test_upload_panel_into_a_buffer_the_loader_cleared():
    """upload_panel(out=..., clear=False): no memset in front of the pull -- a loader that double-buffers its panels
    zeroes a buffer straight after the step that used it.  Unmasked rows arrive, masked rows keep what the buffer
    held (zeros here; the stale value when nobody cleared it)."""
    from deeplearninginassetpricing_paperreplication_b200 import upload_panel
    T, N, F = 9, 700, 46
    rng = np.random.default_rng(6)
    mask = torch.from_numpy(rng.random((T, N)) < 0.4)
    x = torch.from_numpy(rng.standard_normal((T, N, F)).astype(np.float32))
    R = torch.from_numpy((0.1 * rng.standard_normal((T, N))).astype(np.float32)) * mask
    macro = torch.from_numpy(rng.standard_normal((T, 178)).astype(np.float32))
    host = [t.pin_memory() for t in (macro, x, R, mask)]
    for stale in (0.0, 7.0):
        out = [torch.empty_like(t, device=dev()) for t in host]
        out[1].fill_(stale)
        _, d_x, _, d_mask = upload_panel(*host, device=dev(), out=out, clear=False)
        torch.cuda.synchronize()
        xd = d_x.cpu()
        assert d_x is out[1] and torch.equal(d_mask.cpu(), mask)
        assert torch.equal(xd[mask], x[mask])
        assert bool((xd[~mask] == stale).all())


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_graphed_step_replays_the_eager_step_bit_for_bit(precision):
    """GraphedStep (zero_grad + forward + backward in one CUDA graph): every replay reproduces the eager step's
    losses, weights and gradients exactly, also after the parameters changed in place (an optimiser step)."""
    g = load_golden("rand_small")
    params = golden_params(g)
    macro, x, R, mask, _ = _panel(g)
    model = _model_from(params, 178).eval()
    model.precision = precision

    def eager():
        model.zero_grad(set_to_none=True)
        out = model(macro, x, R, mask)
        out["loss"].backward()
        return ([out[k].detach().clone() for k in ("loss", "loss_unconditional", "weights", "portfolio_returns")],
                {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None})

    ref_out, ref_grads = eager()
    gs = model.graphed_step(macro, x, R, mask)
    assert gs.launches == (4 if precision == "bf16x3" else 5)      # fp32: the macro LSTM is a launch of its own
    for trial in range(3):
        if trial == 2:                                   # parameters move in place, like optimizer.step()
            with torch.no_grad():
                for p in model.parameters():
                    p.mul_(1.01)
            ref_out, ref_grads = eager()
        out = gs.step()
        torch.cuda.synchronize()
        for a, key in zip(ref_out, ("loss", "loss_unconditional", "weights", "portfolio_returns")):
            assert torch.equal(a, out[key]), key
        for k, p in model.named_parameters():
            if k in ref_grads:
                assert torch.equal(ref_grads[k], p.grad), k


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_graphed_step_draws_a_fresh_dropout_mask_every_replay(precision):
    """Train mode: the dropout seed of a GraphedStep lives in device memory and the graph advances it.  A replay
    equals the eager step run with that same device seed, and consecutive replays differ."""
    g = load_golden("rand_small")
    params = golden_params(g)
    macro, x, R, mask, _ = _panel(g)
    model = _model_from(params, 178).train()
    model.precision = precision
    model.sdf_net.dropout_p = 0.3
    gs = model.graphed_step(macro, x, R, mask)
    assert gs.seed is not None
    out = gs.step()
    torch.cuda.synchronize()
    loss1 = out["loss"].detach().clone()
    grad1 = model.sdf_net.fc_layers[0].weight.grad.clone()
    model.zero_grad(set_to_none=True)
    model._seed_dev = gs.seed.clone()                    # the seed that replay used
    try:
        ref = model(macro, x, R, mask)
        ref["loss"].backward()
    finally:
        model._seed_dev = None
    assert torch.equal(ref["loss"].detach(), loss1)
    assert torch.equal(model.sdf_net.fc_layers[0].weight.grad, grad1)
    out = gs.step()
    torch.cuda.synchronize()
    assert not torch.equal(out["loss"].detach(), loss1)


# ---- full step against the reference's own outputs ---------------------------------------------------
@pytest.mark.parametrize("phase", PHASES)
def test_step_matches_reference_goldens(golden, phase):
    params = golden_params(golden)
    M = golden["macro"].shape[1]
    model = _model_from(params, M).eval()
    macro, x, R, mask, hidden = _panel(golden)
    frozen = phase == "moment"
    for p in model.sdf_net.parameters():
        p.requires_grad = not frozen
    out = model(macro, x, R, mask, hidden=hidden, phase=phase)
    out["loss"].backward()
    torch.cuda.synchronize()
    for key in ("loss", "loss_unconditional", "loss_conditional"):
        a, b = float(out[key]), float(golden[f"{phase}/{key}"])
        assert abs(a - b) <= LOSS_RTOL * abs(b) + 1e-30, (key, a, b)
    assert rel_err(_np(out["weights"]), golden[f"{phase}/weights"]) < WEIGHT_TOL
    assert np.all(_np(out["weights"])[~golden["mask"]] == 0.0)
    assert rel_err(_np(out["portfolio_returns"]), golden[f"{phase}/portfolio_returns"]) < WEIGHT_TOL
    assert abs(float(out["sharpe"]) - float(golden[f"{phase}/sharpe"])) < 1e-4 * abs(float(golden[f"{phase}/sharpe"])) + 1e-7
    assert np.abs(_np(out["hidden"][0]) - golden[f"{phase}/hidden_h"]).max() < 2e-6
    assert np.abs(_np(out["hidden"][1]) - golden[f"{phase}/hidden_c"]).max() < 2e-6
    assert tuple(out["hidden"][0].shape) == (1, 1, 4)
    seen = 0
    for k, p in model.named_parameters():
        gk = f"{phase}/grad/{k}"
        if gk not in golden:
            assert p.grad is None, k
            continue
        seen += 1
        ref = golden[gk]
        if k == "sdf_net.output_proj.bias":
            continue
        err = np.abs(_np(p.grad) - ref).max() / max(np.abs(ref).max(), 1e-30)
        assert err < 1e-3, (k, err)        # the reference's own fp32 autograd noise is ~1e-4 here
    assert seen >= 2
    if phase == "conditional":
        q = out["moments"]
        assert tuple(q.shape) == golden["moments"].shape
        assert np.abs(_np(q) - golden["moments"]).max() < 2e-6


def test_normalized_weights_match_reference(golden):
    params = golden_params(golden)
    model = _model_from(params, golden["macro"].shape[1]).eval()
    macro, x, R, mask, hidden = _panel(golden)
    with torch.no_grad():
        wn, new_hidden = model.get_weights(macro, x, mask, hidden=hidden, normalized=True)
        w, _ = model.get_weights(macro, x, mask, hidden=hidden)
        w2, _ = model.sdf_net(macro, x, mask, hidden)
        f = model.get_sdf_factor(macro, x, R, mask, hidden=hidden)
    assert rel_err(_np(wn), golden["normalized_weights"]) < WEIGHT_TOL
    assert rel_err(_np(w), golden["conditional/weights"]) < WEIGHT_TOL
    assert torch.equal(w, w2)
    assert rel_err(_np(f), golden["sdf_factor"]) < WEIGHT_TOL
    live = golden["mask"].any(axis=1)
    s = (np.abs(_np(wn)) * golden["mask"]).sum(axis=1)
    assert np.allclose(s[live & (np.abs(golden["conditional/weights"]).sum(axis=1) > 1e-6)], 1.0, atol=1e-5)


# ---- oracle parity on shapes the goldens do not cover ---------------------------------------------------
@pytest.mark.parametrize("T,N,M", [(1, 1, 8), (3, 127, 8), (5, 129, 178), (7, 1000, 178), (2, 257, 3),
                                   (16, 513, 178)])
@pytest.mark.parametrize("phase", PHASES)
def test_step_vs_oracle_shapes(T, N, M, phase):
    macro, x, R, mask = make_panel(T, N, M, seed=T * 1000 + N)
    if T > 2:
        mask[1] = False                 # an empty month
    if N > 3:
        mask[:, 2] = False              # a never-valid stock
    x = x * mask[:, :, None]; R = R * mask
    torch.manual_seed(N)
    model = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": M, "individual_feature_dim": 46}).to(dev()).eval()
    _check_against_cf(model, macro, x, R, mask, phase)


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_series_longer_than_256_months(precision):
    """T = 300: the month-summed gradients walk the series in sweeps of 256 months (macro_lstm.cu, wgrad_tasks: the
    first sweep's operands are prefetched, later ones loaded on the spot), the BPTT scan's chunks are 19 months long,
    and the LSTM publishes 19 groups.  fwd + bwd against the fp64 closed form; the frozen-SDF phase takes the
    stand-alone month-summed-gradient kernel over the same series."""
    T, N, M = 300, 70, 8
    macro, x, R, mask = make_panel(T, N, M, seed=300)
    mask[7] = False
    x = x * mask[:, :, None]; R = R * mask
    torch.manual_seed(5)
    model = AssetPricingGAN({"b200_precision": precision, "macro_feature_dim": M, "individual_feature_dim": 46}).to(dev()).eval()
    tol = GRAD_TOL if precision == "fp32" else 4e-2
    for phase in ("conditional", "moment"):
        _check_against_cf(model, macro, x, R, mask, phase, grad_tol=tol)


def test_no_macro_and_other_feature_counts():
    for F in (46, 13, 64):
        _, x, R, mask = make_panel(6, 200, 0, F=F, seed=F)
        torch.manual_seed(1)
        model = AssetPricingGAN({"b200_precision": "fp32", "individual_feature_dim": F}).to(dev()).eval()
        _check_against_cf(model, None, x, R, mask, "conditional")
        macro, x, R, mask = make_panel(6, 200, 20, F=F, seed=F)
        model = AssetPricingGAN({"b200_precision": "fp32", "individual_feature_dim": F, "macro_feature_dim": 20,
                                 "num_units_rnn": [6]}).to(dev()).eval()
        _check_against_cf(model, macro, x, R, mask, "conditional")


def test_unaligned_panel_view():
    """x as a view whose first element is only 8-byte aligned, odd N: exercises the copy shift."""
    macro, x, R, mask = make_panel(5, 131, 8, seed=9)
    d = dev()
    big = torch.zeros(5 * 131 * 46 + 2, device=d)
    xv = big[2:].view(5, 131, 46)
    xv.copy_(x)
    assert xv.data_ptr() % 16 == 8
    torch.manual_seed(2)
    model = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": 8, "individual_feature_dim": 46}).to(d).eval()
    o1 = model(macro.to(d), xv, R.to(d), mask.to(d))
    o2 = model(macro.to(d), x.to(d), R.to(d), mask.to(d))
    assert torch.equal(o1["weights"], o2["weights"]) and torch.equal(o1["loss"], o2["loss"])


def test_all_masked_panel():
    macro, x, R, mask = make_panel(4, 150, 8, seed=1)
    mask[:] = False
    x = x * 0; R = R * 0
    model = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": 8, "individual_feature_dim": 46}).to(dev()).eval()
    out = model(macro.to(dev()), x.to(dev()), R.to(dev()), mask.to(dev()))
    out["loss"].backward()
    assert float(out["loss"]) == 0.0 and float(out["weights"].abs().max()) == 0.0
    assert all(torch.isfinite(p.grad).all() for p in model.parameters() if p.grad is not None)


def test_bit_reproducible():
    macro, x, R, mask = make_panel(12, 700, 178, seed=4)
    d = dev()
    torch.manual_seed(3)
    model = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": 178, "individual_feature_dim": 46}).to(d).eval()
    args = (macro.to(d), x.to(d), R.to(d), mask.to(d))
    res = []
    for _ in range(2):
        model.zero_grad(set_to_none=True)
        out = model(*args)
        out["loss"].backward()
        res.append((out["loss"].clone(), out["weights"].clone(), [p.grad.clone() for p in model.parameters()]))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
    assert all(torch.equal(a, b) for a, b in zip(res[0][2], res[1][2]))


def test_grad_accumulates_and_cotangents():
    """loss.backward() accumulates into .grad like autograd; gradients are linear in the cotangent."""
    macro, x, R, mask = make_panel(6, 300, 8, seed=5)
    d = dev()
    torch.manual_seed(3)
    model = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": 8, "individual_feature_dim": 46}).to(d).eval()
    args = (macro.to(d), x.to(d), R.to(d), mask.to(d))
    model(*args)["loss"].backward()
    g1 = [p.grad.clone() for p in model.parameters()]
    (2.0 * model(*args)["loss"]).backward()
    for p, a in zip(model.parameters(), g1):
        assert torch.allclose(p.grad, 3.0 * a, rtol=1e-5, atol=1e-12)
    # loss_unconditional and portfolio_returns are differentiable outputs too
    model.zero_grad(set_to_none=True)
    out = model(*args)
    (out["loss_unconditional"] + out["portfolio_returns"].sum() * 1e-3).backward()
    assert model.sdf_net.output_proj.weight.grad.abs().max() > 0


def test_trajectory_matches_oracle():
    """Five Adam steps of phase 1 then five of phase 3 (dropout 0): losses track the CPU oracle."""
    from oracle import torch_port as tp
    macro, x, R, mask = make_panel(24, 400, 8, seed=6)
    d = dev()
    torch.manual_seed(5)
    model = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.0}).to(d)
    ref = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in model.state_dict().items()}
    opt = torch.optim.Adam(model.sdf_net.parameters(), lr=1e-3)
    ropt = torch.optim.Adam([ref[k] for k in tp.SDF_KEYS], lr=1e-3)
    args = (macro.to(d), x.to(d), R.to(d), mask.to(d))
    for phase in ["unconditional"] * 5 + ["conditional"] * 5:
        model.train()
        opt.zero_grad(); ropt.zero_grad()
        out = model(*args, phase=phase)
        out["loss"].backward()
        torch.nn.utils.clip_grad_norm_(model.sdf_net.parameters(), 1.0)
        opt.step()
        rout = tp.forward(ref, macro, x, R, mask, phase=phase)
        rout["loss"].backward()
        torch.nn.utils.clip_grad_norm_([ref[k] for k in tp.SDF_KEYS], 1.0)
        ropt.step()
        a, b = float(out["loss"]), float(rout["loss"])
        assert abs(a - b) <= 1e-4 * abs(b), (phase, a, b)


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_three_phase_schedule_matches_oracle(precision):
    """The 3-phase loop of src/train.py:238-394 in miniature (two Adam optimisers, clip at 1.0, SDF net
    frozen with requires_grad=False in phase 2 as at :309-312, dropout 0), driven through the drop-in
    classes on the GPU and through the oracle port on the CPU: the loss trajectories and the final
    normalised-weight Sharpe ratio (evaluate(), src/train.py:133-137) must agree."""
    from oracle import torch_port as tp
    macro, x, R, mask = make_panel(30, 500, 8, seed=16)
    d = dev()
    torch.manual_seed(8)
    model = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.0,
                             "b200_precision": precision}).to(d)
    ref = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in model.state_dict().items()}
    opt_s = torch.optim.Adam(model.sdf_net.parameters(), lr=1e-3)
    opt_m = torch.optim.Adam(model.moment_net.parameters(), lr=1e-3)
    ropt_s = torch.optim.Adam([ref[k] for k in tp.SDF_KEYS], lr=1e-3)
    ropt_m = torch.optim.Adam([ref[k] for k in tp.MOMENT_KEYS], lr=1e-3)
    args = (macro.to(d), x.to(d), R.to(d), mask.to(d))
    tol = 2e-4 if precision == "fp32" else 1e-3
    schedule = [("unconditional", "sdf")] * 4 + [("moment", "moment")] * 3 + [("conditional", "sdf")] * 4
    for phase, scope in schedule:
        model.train()
        frozen = phase == "moment"
        for p in model.sdf_net.parameters():
            p.requires_grad = not frozen
        for k in tp.SDF_KEYS:
            ref[k].requires_grad_(not frozen)
        opt, ropt = (opt_s, ropt_s) if scope == "sdf" else (opt_m, ropt_m)
        group = model.sdf_net.parameters() if scope == "sdf" else model.moment_net.parameters()
        rgroup = [ref[k] for k in (tp.SDF_KEYS if scope == "sdf" else tp.MOMENT_KEYS)]
        opt.zero_grad(); ropt.zero_grad()
        out = model(*args, phase=phase)
        out["loss"].backward()
        torch.nn.utils.clip_grad_norm_(group, 1.0)
        opt.step()
        rout = tp.forward(ref, macro, x, R, mask, phase=phase)
        rout["loss"].backward()
        torch.nn.utils.clip_grad_norm_(rgroup, 1.0)
        ropt.step()
        a, b = float(out["loss"]), float(rout["loss"])
        assert abs(a - b) <= tol * abs(b), (phase, a, b)
    model.eval()
    with torch.no_grad():
        wn, _ = model.get_weights(args[0], args[1], args[3], normalized=True)
        pr = (wn * args[2] * args[3].float()).sum(dim=1).cpu()
        rwn, _ = tp.normalized_weights({k: v.detach() for k, v in ref.items()}, macro, x, mask)
        rpr = (rwn * R * mask.float()).sum(dim=1)
    sa, sb = float(pr.mean() / pr.std()), float(rpr.mean() / rpr.std())
    assert abs(sa - sb) <= 5e-3 * abs(sb) + 1e-4, (sa, sb)


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_dropout_train_mode(precision):
    macro, x, R, mask = make_panel(8, 600, 8, seed=7)
    d = dev()
    torch.manual_seed(5)
    model = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.05,
                             "b200_precision": precision}).to(d)
    args = (macro.to(d), x.to(d), R.to(d), mask.to(d))
    model.eval()
    base = model(*args)
    model.train()
    torch.manual_seed(11); o1 = model(*args)
    torch.manual_seed(11); o2 = model(*args)
    torch.manual_seed(12); o3 = model(*args)
    assert torch.equal(o1["weights"], o2["weights"])              # same seed -> same mask
    assert not torch.equal(o1["weights"], o3["weights"])
    assert not torch.equal(o1["weights"], base["weights"])
    rel = float((o1["weights"] - base["weights"]).abs().max() / base["weights"].abs().max())
    assert 1e-3 < rel < 2.0
    # forward/backward use the same mask: directional derivative along W3 matches the gradient
    w3 = model.sdf_net.output_proj.weight
    torch.manual_seed(11)
    model.zero_grad(set_to_none=True)
    l0 = model(*args, phase="unconditional")["loss"]; l0.backward()
    gdir = w3.grad.clone()
    direction = torch.randn_like(w3)
    eps = 1e-2
    with torch.no_grad():
        w3.add_(eps * direction)
    torch.manual_seed(11); lp = model(*args, phase="unconditional")["loss"]
    with torch.no_grad():
        w3.add_(-2 * eps * direction)
    torch.manual_seed(11); lm = model(*args, phase="unconditional")["loss"]
    fd = float(lp - lm) / (2 * eps)
    an = float((gdir * direction).sum())
    assert abs(fd - an) <= 5e-2 * abs(an) + 1e-9, (fd, an)


def test_dropout_mask_is_the_same_in_both_precisions():
    """Counter-based dropout is keyed on (seed, stock-month, layer, unit) only: the tensor-core kernels
    drop exactly the units the fp32 kernels drop, so a training run does not depend on the arithmetic."""
    macro, x, R, mask = make_panel(6, 900, 8, seed=17)
    d = dev()
    args = (macro.to(d), x.to(d), R.to(d), mask.to(d))
    torch.manual_seed(5)
    m32 = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.05}).to(d)
    mtc = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.05,
                           "b200_precision": "bf16x3"}).to(d)
    mtc.load_state_dict(m32.state_dict())
    m32.train(); mtc.train()
    torch.manual_seed(3); a = m32(*args); a["loss"].backward()
    torch.manual_seed(3); b = mtc(*args); b["loss"].backward()
    assert rel_err(_np(b["weights"]), _np(a["weights"])) < 5e-5
    assert abs(float(a["loss"]) - float(b["loss"])) <= 1e-5 * abs(float(a["loss"]))
    ga, gb = m32.sdf_net.fc_layers[3].weight.grad, mtc.sdf_net.fc_layers[3].weight.grad
    assert float((ga - gb).norm() / ga.norm()) < 2e-2      # same mask; residual = ReLU-kink noise (see test_tc_parity)


def test_errors_are_loud():
    d = dev()
    macro, x, R, mask = make_panel(4, 50, 8, seed=8)
    model = AssetPricingGAN({"b200_precision": "fp32", "macro_feature_dim": 8, "individual_feature_dim": 46}).to(d)
    with pytest.raises(ValueError):
        model(macro.to(d), x.to(d).double(), R.to(d), mask.to(d))
    with pytest.raises(ValueError):
        model(macro.to(d), x.to(d).transpose(0, 1), R.to(d), mask.to(d))
    for cfg in ({"hidden_dim": [64, 64, 32]}, {"residual_loss_factor": 0.1, "normalize_w": False},
                {"num_units_rnn": [4, 4]}):
        bad = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, **cfg}).to(d)
        with pytest.raises(NotImplementedError):
            bad(macro.to(d), x.to(d), R.to(d), mask.to(d))
