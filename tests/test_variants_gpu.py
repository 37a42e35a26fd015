"""the reference's other configurations on the drop-in, against goldens written by the reference
itself (tests/golden/make_golden.py `variants`): the notebooks' hidden_dim=[64,32] + use_rnn=False +
hidden_dim_moment=[32] (notebooks/demo.ipynb, demo_synthetic.ipynb), each of those alone, normalize_w=False, and
narrow hidden layers with weighted_loss=False.  All three phases, both arithmetic paths, losses / weights /
portfolio returns / normalised weights / every gradient."""
import json

import numpy as np
import pytest
import torch

from _dlap_testlib import PHASES, golden_params, load_golden, rel_err
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
from test_parity_gpu import LOSS_RTOL, WEIGHT_TOL, _np, dev

pytestmark = pytest.mark.gpu
VARIANTS = ["v_notebook", "v_6432", "v_norm_off", "v_no_rnn", "v_mom_hidden", "v_narrow", "v_residual"]


def _model(g, precision):
    cfg = json.loads(str(g["cfg_json"]))
    model = AssetPricingGAN({"macro_feature_dim": int(g["macro"].shape[1]), "individual_feature_dim": 46,
                             "b200_precision": precision, **cfg})
    model.load_state_dict({k: torch.as_tensor(v) for k, v in golden_params(g).items()})     # same keys, same shapes
    return model.to(dev()).eval(), cfg


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
@pytest.mark.parametrize("phase", PHASES)
@pytest.mark.parametrize("name", VARIANTS)
def test_variant_step_matches_the_reference(name, phase, precision):
    g = load_golden(name)
    model, cfg = _model(g, precision)
    d = dev()
    macro, x, R, mask = [torch.from_numpy(g[k]).to(d) for k in ("macro", "x", "R", "mask")]
    for p in model.sdf_net.parameters():
        p.requires_grad = phase != "moment"
    out = model(macro, x, R, mask, phase=phase)
    out["loss"].backward()
    torch.cuda.synchronize()
    for key in ("loss", "loss_unconditional", "loss_conditional", "loss_residual"):
        if f"{phase}/{key}" not in g:                        # goldens older than the residual loss: it is zero there
            assert float(out[key]) == 0.0
            continue
        a, b = float(out[key]), float(g[f"{phase}/{key}"])
        assert abs(a - b) <= LOSS_RTOL * abs(b) + 1e-30, (key, a, b)
    w = _np(out["weights"])
    assert np.all(w[~g["mask"]] == 0.0)
    assert rel_err(w, g[f"{phase}/weights"]) < WEIGHT_TOL
    assert rel_err(_np(out["portfolio_returns"]), g[f"{phase}/portfolio_returns"]) < WEIGHT_TOL
    # (a Sharpe ratio near zero is a cancelling mean over ten months: absolute floor)
    assert abs(float(out["sharpe"]) - float(g[f"{phase}/sharpe"])) <= 1e-4 * abs(float(g[f"{phase}/sharpe"])) + 1e-5
    if f"{phase}/hidden_h" in g:
        assert np.abs(_np(out["hidden"][0]) - g[f"{phase}/hidden_h"]).max() < 2e-6
    else:
        assert out["hidden"] is None
    tol_max, tol_norm = (2e-4, 2e-4) if precision == "fp32" else (4e-2, 1e-2)
    seen = 0
    for k, p in model.named_parameters():
        gk = f"{phase}/grad/{k}"
        if gk not in g:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
            continue
        assert p.grad is not None, k
        ref = g[gk]
        assert tuple(p.grad.shape) == ref.shape, k
        if k == "sdf_net.output_proj.bias" and cfg.get("normalize_w", True):
            assert float(p.grad.abs().max()) < 1e-9          # identically zero with demeaning
            continue
        got = _np(p.grad)
        emax = np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30)
        enorm = np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-30)
        assert emax < tol_max and enorm < tol_norm, (k, emax, enorm)
        seen += 1
    assert seen >= 2


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
@pytest.mark.parametrize("name", VARIANTS)
def test_variant_inference_matches_the_reference(name, precision):
    g = load_golden(name)
    model, _ = _model(g, precision)
    d = dev()
    macro, x, R, mask = [torch.from_numpy(g[k]).to(d) for k in ("macro", "x", "R", "mask")]
    with torch.no_grad():
        wn, _ = model.get_weights(macro, x, mask, normalized=True)
        out = model(macro, x, R, mask, phase="conditional")              # evaluate(): the same panel right after
        f = model.get_sdf_factor(macro, x, R, mask)
    assert rel_err(_np(wn), g["normalized_weights"]) < WEIGHT_TOL
    assert rel_err(_np(f), g["sdf_factor"]) < WEIGHT_TOL
    assert abs(float(out["loss"]) - float(g["conditional/loss"])) <= LOSS_RTOL * abs(float(g["conditional/loss"]))
    assert rel_err(_np(out["moments"]), g["moments"]) < 1e-4


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_residual_loss_in_a_graphed_epoch(precision):
    """residual_loss_factor > 0 through GraphedStep + FusedClipAdam: the [T]-sized torch ops of the regulariser are part
    of the capture; replays follow the eager model stepped by torch.optim.Adam, and the report carries the loss."""
    from deeplearninginassetpricing_paperreplication_b200 import FusedClipAdam
    g = load_golden("v_residual")
    a, _ = _model(g, precision)
    b, _ = _model(g, precision)
    d = dev()
    macro, x, R, mask = [torch.from_numpy(g[k]).to(d) for k in ("macro", "x", "R", "mask")]
    ropt = torch.optim.Adam(a.sdf_net.parameters(), lr=1e-3)
    opt = FusedClipAdam(b.sdf_net.parameters(), lr=1e-3)
    gs = b.graphed_step(macro, x, R, mask, phase="conditional", optimizer=opt, max_norm=1.0)
    for epoch in range(5):
        ropt.zero_grad()
        out = a(macro, x, R, mask, phase="conditional")
        out["loss"].backward()
        torch.nn.utils.clip_grad_norm_(a.sdf_net.parameters(), 1.0)
        ropt.step()
        got = gs.step()
        torch.cuda.synchronize()
        ref = float(out["loss"].detach())
        assert abs(float(got["loss"].detach()) - ref) <= 2e-3 * abs(ref), epoch
        rep = gs.report.cpu().numpy()
        assert rep[4] == float(got["loss_residual"]) and rep[4] > 0.5
        del out


def test_still_unsupported_configurations_raise():
    d = dev()
    g = load_golden("v_6432")
    macro, x, R, mask = [torch.from_numpy(g[k]).to(d) for k in ("macro", "x", "R", "mask")]
    for cfg in ({"hidden_dim": [64, 64, 64]}, {"hidden_dim": [128, 64]}, {"num_units_rnn": [4, 4]},
                {"num_condition_moment": 4}, {"residual_loss_factor": 0.1, "normalize_w": False}):
        bad = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, **cfg}).to(d)
        with pytest.raises(NotImplementedError):
            bad(macro, x, R, mask)
