"""The oracle against the reference's own outputs (tests/golden/*.npz, written by
tests/golden/make_golden.py from /root/reference).  CPU only.

torch_port runs the same torch ops in the same order as the reference, so on the host that wrote
the fixtures it reproduces them bit for bit.  torch's CPU GEMM/LSTM kernels (oneDNN, MKL) dispatch
on the host ISA (AVX2, AVX-512, AMX) and sum in a different order on another CPU model, so on any
other host the comparison is "equal, or within HOST_RTOL of the tensor's largest entry" -- a few
fp32 ulps, well inside the path's 1e-5 loss tolerance.  Integer/index/mask fixtures and the initial
parameters stay bit-exact everywhere.  closed_form (fp64, hand-derived backward) must agree to
fp32 rounding.
"""
import numpy as np
import pytest
import torch

from oracle import closed_form as cf
from oracle import torch_port as tp
from _dlap_testlib import load_golden

PHASES = ("unconditional", "moment", "conditional")
HOST_RTOL = 4e-6          # measured between an AVX2 and an AVX-512/AMX host: <= 1.6e-6


def _same(a, b):
    """Bit-equal, or equal up to the summation order of the host's BLAS kernels (see header)."""
    a = np.asarray(a); b = np.asarray(b)
    if a.shape != b.shape or a.dtype != b.dtype:
        return False
    if np.array_equal(a, b):
        return True
    a = a.astype(np.float64); b = b.astype(np.float64)
    return bool(np.abs(a - b).max() <= HOST_RTOL * max(np.abs(b).max(), 1e-30))


def _inputs(g):
    params = {k[len("param/"):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("param/")}
    hidden = None
    if "h0" in g:
        hidden = (torch.from_numpy(g["h0"]), torch.from_numpy(g["c0"]))
    return (params, torch.from_numpy(g["macro"]), torch.from_numpy(g["x"]),
            torch.from_numpy(g["R"]), torch.from_numpy(g["mask"]), hidden)


def test_init_matches_reference(golden):
    """Same seed -> same initial parameters as AssetPricingGAN(cfg) (src/model.py:298-340)."""
    p = tp.init_params(golden["macro"].shape[1], golden["x"].shape[2], int(golden["seed"]))
    for k in tp.PARAM_KEYS:
        assert np.array_equal(p[k].numpy(), golden["param/" + k]), k


@pytest.mark.parametrize("phase", PHASES)
def test_torch_port_matches_reference(golden, phase):
    torch.set_num_threads(1)
    params, macro, x, R, mask, hidden = _inputs(golden)
    keys = tp.MOMENT_KEYS if phase == "moment" else tp.PARAM_KEYS
    out, grads = tp.step(params, macro, x, R, mask, phase=phase, hidden=hidden, grad_keys=keys)
    for key in ("weights", "loss", "loss_unconditional", "loss_conditional", "sharpe",
                "portfolio_returns"):
        assert _same(out[key].numpy(), golden[f"{phase}/{key}"]), key
    assert _same(out["hidden"][0].numpy(), golden[f"{phase}/hidden_h"])
    assert _same(out["hidden"][1].numpy(), golden[f"{phase}/hidden_c"])
    seen = 0
    for k in keys:
        gk = f"{phase}/grad/{k}"
        if gk in golden:
            seen += 1
            assert grads[k] is not None, k
            if k == "sdf_net.output_proj.bias":
                # identically zero in exact arithmetic (demeaning): both sides hold fp32 noise only
                assert np.abs(grads[k].numpy()).max() < 1e-6 and np.abs(golden[gk]).max() < 1e-6
                continue
            assert _same(grads[k].numpy(), golden[gk]), k
        else:  # phase 1: the moment net receives no gradient at all (src/model.py:524-528)
            assert grads[k] is None, k
    assert seen >= 2


@pytest.mark.parametrize("phase", PHASES)
def test_torch_port_residual_loss_matches_reference(phase):
    """residual_loss_factor = 0.25 (src/model.py:435-483, :543-548) on a panel with an empty month and a month with a
    single listed stock: the port against the reference's own numbers, gradients included."""
    torch.set_num_threads(1)
    g = load_golden("v_residual")
    params, macro, x, R, mask, hidden = _inputs(g)
    keys = tp.MOMENT_KEYS if phase == "moment" else tp.PARAM_KEYS
    out, grads = tp.step(params, macro, x, R, mask, phase=phase, hidden=hidden, grad_keys=keys,
                         residual_loss_factor=0.25)
    for key in ("weights", "loss", "loss_unconditional", "loss_conditional", "loss_residual", "portfolio_returns"):
        assert _same(out[key].numpy(), g[f"{phase}/{key}"]), key
    assert float(out["loss_residual"]) > 0.5
    for k in keys:
        if f"{phase}/grad/{k}" not in g:                    # phase 1 never touches the moment network
            assert grads[k] is None, k
            continue
        assert _same(grads[k].numpy(), g[f"{phase}/grad/{k}"]), k


def test_residual_terms_value_and_gradient_on_cpu():
    """functional.residual_terms (the host-side part of residual_loss_factor > 0: [T]-sized torch ops, device-agnostic)
    against the port of compute_residual_loss: the value, and c_t (R - beta_t w) against autograd's dL/dw -- an empty
    month, a month with one listed stock (skipped), a month whose weights vanish (denominator only)."""
    from deeplearninginassetpricing_paperreplication_b200 import functional as Fn
    torch.manual_seed(0)
    T, N = 9, 40
    mask = torch.rand(T, N) < 0.5
    mask[2] = False
    mask[4] = False; mask[4, 3] = True
    w = torch.randn(T, N) * 0.01 * mask
    w[6] = 0.0
    w.requires_grad_(True)
    R = torch.randn(T, N) * 0.1 * mask
    loss = tp.residual_loss(w, R, mask)
    loss.backward()
    got, c, beta = Fn.residual_terms(w.detach(), R, mask)
    assert abs(float(got) - float(loss.detach())) <= 1e-6 * float(loss.detach())
    g = (c[:, None] * (R.double() - beta[:, None] * w.detach().double())) * mask
    assert float((g - w.grad.double()).abs().max()) <= 1e-6 * float(w.grad.abs().max())
    assert float(c[2]) == 0.0 and float(c[4]) == 0.0 and float(c[6]) == 0.0
    none, c0, _ = Fn.residual_terms(torch.zeros(T, N), R, mask)          # no month with weights: the loss is 0
    assert float(none) == 0.0 and not c0.any()


def test_torch_port_moments_and_normalized(golden):
    torch.set_num_threads(1)
    params, macro, x, R, mask, hidden = _inputs(golden)
    q = tp.moment_conditions(params, macro, x)
    assert q.shape == golden["moments"].shape and not q.is_contiguous()
    assert _same(q.contiguous().numpy(), golden["moments"])
    # evaluate() runs under @torch.no_grad (src/train.py:106); oneDNN's LSTM takes a different
    # code path there and differs from grad mode in the last bit, so mirror it.
    with torch.no_grad():
        wn, _ = tp.normalized_weights(params, macro, x, mask, hidden)
    assert _same(wn.numpy(), golden["normalized_weights"])


def _rel(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


@pytest.mark.parametrize("phase", PHASES)
def test_closed_form_matches_reference(golden, phase):
    params = {k[len("param/"):]: v for k, v in golden.items() if k.startswith("param/")}
    h0 = golden.get("h0"); c0 = golden.get("c0")
    out = cf.forward(params, golden["macro"], golden["x"], golden["R"], golden["mask"],
                     h0=h0, c0=c0, phase=phase)
    # fixtures are fp32 results of the reference; 2e-5 relative covers its rounding
    assert _rel(out["weights"], golden[f"{phase}/weights"]) < 2e-5
    assert _rel(out["portfolio_returns"], golden[f"{phase}/portfolio_returns"]) < 2e-5
    for key in ("loss", "loss_unconditional", "loss_conditional", "sharpe"):
        ref = float(golden[f"{phase}/{key}"])
        assert abs(float(out[key]) - ref) <= 2e-5 * abs(ref) + 1e-12, key
    # masked cells are exactly zero, per-month mean over valid cells is zero
    assert np.all(out["weights"][~golden["mask"]] == 0.0)
    grads = cf.backward(params, golden["macro"], golden["x"], golden["R"], golden["mask"], out,
                        phase=phase)
    for k in cf.PARAM_KEYS:
        gk = f"{phase}/grad/{k}"
        if gk not in golden:
            continue
        ref = golden[gk]
        scale = max(np.abs(ref).max(), 1e-12)
        err = np.abs(grads[k].reshape(ref.shape) - ref).max()
        if k.endswith("output_proj.bias") and k.startswith("sdf_net"):
            # identically zero in exact arithmetic (demeaning); the reference leaves fp32 noise
            assert abs(grads[k]).max() < 1e-12 and np.abs(ref).max() < 1e-6
            continue
        assert err / scale < 5e-4, (k, err / scale)


def test_closed_form_normalized_weights(golden):
    params = {k[len("param/"):]: v for k, v in golden.items() if k.startswith("param/")}
    out = cf.forward(params, golden["macro"], golden["x"], golden["R"], golden["mask"],
                     h0=golden.get("h0"), c0=golden.get("c0"))
    wn = cf.normalized_weights(out["weights"], golden["mask"])
    assert _rel(wn, golden["normalized_weights"]) < 2e-5


def test_mask_contract_bit_exact():
    """src/data_loader.py:50-65 -- sentinel, NaN and threshold handling."""
    g = load_golden("mask_raw")
    raw = g["raw"]
    mask, ret, feat = cf.mask_from_raw(raw[:, :, 0], raw[:, :, 1:])
    assert np.array_equal(mask, g["mask"])
    assert np.array_equal(ret, g["returns"])
    assert np.array_equal(feat, g["features"])
    assert not mask[2, 3] and not mask[4, 7] and not mask[5, 1]


def test_loader_contract_bit_exact():
    """src/data_loader.py:42-99, :212-237 on raw files written by make_golden.py `loader`: mask, zero-fill (fp32 and
    fp64 files), macro statistics and z-scores in numpy's summation order, validation file with training statistics."""
    g = load_golden("loader")
    idx = g["macro_idx"]
    stats = None
    for split in ("train", "valid"):
        raw = g[f"{split}/raw"]
        mask, ret, feat = cf.mask_from_raw(raw[:, :, 0], raw[:, :, 1:])
        assert np.array_equal(mask, g[f"{split}/mask"])
        assert ret.tobytes() == g[f"{split}/returns"].tobytes()            # bytes: -0.0 and +0.0 differ
        assert feat.tobytes() == g[f"{split}/features"].tobytes()
        z, mean, std = cf.macro_zscore(g[f"{split}/macro_raw"][:, idx], *(stats or (None, None)), pairwise=True)
        assert z.tobytes() == g[f"{split}/macro"].tobytes()
        assert mean.tobytes() == g[f"{split}/mean"].tobytes() and std.tobytes() == g[f"{split}/std"].tobytes()
        assert np.array_equal(mask.sum(axis=1).astype(np.float32), g[f"{split}/date_count"])
        stats = stats or (mean, std)
    z, mean, std = cf.macro_zscore(g["train/macro_raw"])                      # no macro_idx: row-after-row sums
    assert z.tobytes() == g["train/macro_all"].tobytes() and std.tobytes() == g["train/std_all"].tobytes()
    z, mean, std = cf.macro_zscore(g["long/macro_raw"][:, g["long/idx"]], pairwise=True)      # 300 months: recursive split
    assert z.tobytes() == g["long/macro"].tobytes() and mean.tobytes() == g["long/mean"].tobytes()
    assert std.tobytes() == g["long/std"].tobytes()
    z, mean, std = cf.macro_zscore(g["long/macro_raw"])
    assert z.tobytes() == g["long/macro_all"].tobytes() and mean.tobytes() == g["long/mean_all"].tobytes()
    assert not g["train/mask"][5].any() and g["train/mask"].any()
    raw64 = g["f64/raw"]
    mask, ret, feat = cf.mask_from_raw(raw64[:, :, 0].astype(np.float32), raw64[:, :, 1:].astype(np.float32))
    assert np.array_equal(mask, g["f64/mask"]) and ret.tobytes() == g["f64/returns"].tobytes()
    assert feat.tobytes() == g["f64/features"].tobytes()
    T, cols = cf.small_sample(g["train/mask"], 12, 40)
    assert np.array_equal(g["train/mask"][:T][:, cols], g["small/mask"])
    assert np.array_equal(g["train/features"][:T][:, cols], g["small/individual_features"])

