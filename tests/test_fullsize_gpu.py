"""Size-independent properties at BASELINE.json's full shapes (the oracle is too slow / too large there):
exact zeros at masked cells, zero cross-sectional mean of the weights, |w| normalisation, run-to-run
determinism, agreement of the two arithmetic paths, and cross-checks between outputs
(portfolio return and Sharpe recomputed from the weights with plain torch ops)."""
import pytest
import torch

from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
from test_parity_gpu import dev, make_panel

pytestmark = pytest.mark.gpu


def _paper_model(precision, seed=42):
    torch.manual_seed(seed)
    return AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                            "b200_precision": precision}).to(dev())


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_paper_shape_step_properties(precision):
    d = dev()
    macro, x, R, mask = [t.to(d) for t in make_panel(240, 10_000, 178, seed=1234)]
    model = _paper_model(precision)
    out = model(macro, x, R, mask, phase="conditional")
    out["loss"].backward()
    w, m = out["weights"], mask.float()
    assert torch.isfinite(out["loss"]) and torch.isfinite(out["sharpe"])
    assert float(w[~mask].abs().max()) == 0.0                                   # masked cells: exactly zero
    n_t = m.sum(dim=1).clamp(min=1)
    assert float(((w * m).sum(dim=1) / n_t).abs().max()) < 2e-7                # src/model.py:276-279
    # portfolio return and Sharpe recomputed from the weights with torch ops (src/model.py:363-367, :551)
    pr = (w * R * m).sum(dim=1) / n_t * n_t.mean()
    assert torch.allclose(pr, out["portfolio_returns"], rtol=2e-4, atol=1e-7)
    sharpe = pr.mean() / (pr.std() + 1e-8)
    assert abs(float(sharpe) - float(out["sharpe"])) <= 2e-4 * abs(float(sharpe)) + 1e-6
    # losses recomputed from weights / moments with torch ops (src/model.py:376-385, :418-431)
    sdf = 1 + pr[:, None]
    t_i = m.sum(dim=0).clamp(min=1)
    l_u = (((R * m * sdf).sum(dim=0) / t_i) ** 2).mean()
    q = out["moments"]
    l_c = torch.stack([(((q[k] * R * m * sdf).sum(dim=0) / t_i) ** 2).mean() for k in range(8)]).mean()
    assert abs(float(l_u) - float(out["loss_unconditional"])) <= 1e-4 * float(l_u)
    assert abs(float(l_c) - float(out["loss_conditional"])) <= 1e-4 * float(l_c)
    # gradients: finite, output bias exactly zero (demeaning), deterministic
    g1 = [p.grad.clone() for p in model.parameters()]
    assert all(torch.isfinite(g).all() for g in g1)
    assert float(model.sdf_net.output_proj.bias.grad.abs().max()) == 0.0
    model.zero_grad(set_to_none=True)
    out2 = model(macro, x, R, mask, phase="conditional")
    out2["loss"].backward()
    assert torch.equal(out2["weights"], w) and torch.equal(out2["loss"], out["loss"])
    assert all(torch.equal(a, p.grad) for a, p in zip(g1, model.parameters()))


def test_paper_shape_paths_agree():
    d = dev()
    macro, x, R, mask = [t.to(d) for t in make_panel(240, 10_000, 178, seed=99)]
    a, b = _paper_model("fp32"), _paper_model("bf16x3")
    b.load_state_dict(a.state_dict())
    oa, ob = a(macro, x, R, mask), b(macro, x, R, mask)
    oa["loss"].backward(); ob["loss"].backward()
    assert abs(float(oa["loss"]) - float(ob["loss"])) <= 1e-5 * abs(float(oa["loss"]))
    assert abs(float(oa["loss_unconditional"]) - float(ob["loss_unconditional"])) <= 1e-5 * abs(float(oa["loss_unconditional"]))
    assert float((oa["weights"] - ob["weights"]).abs().max()) <= 1e-4 * float(oa["weights"].abs().max())
    for (k, p), q in zip(a.named_parameters(), b.parameters()):
        if k == "sdf_net.output_proj.bias":
            continue
        assert float((p.grad - q.grad).norm() / p.grad.norm()) < 5e-3, k


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_full_sample_inference_shape(precision):
    """configs[2]: ~600 months, inference only: get_weights(normalized=True) and the SDF factor."""
    d = dev()
    macro, x, R, mask = [t.to(d) for t in make_panel(600, 10_000, 178, seed=7)]
    model = _paper_model(precision).eval()
    with torch.no_grad():
        wn, hidden = model.get_weights(macro, x, mask, normalized=True)
        f = model.get_sdf_factor(macro, x, R, mask)
    m = mask.float()
    assert float(wn[~mask].abs().max()) == 0.0
    s = (wn.abs() * m).sum(dim=1)
    assert torch.allclose(s, torch.ones_like(s), atol=2e-5)                     # src/model.py:586-590
    assert torch.allclose(f, (wn * R * m).sum(dim=1), rtol=1e-5, atol=1e-8) and tuple(hidden[0].shape) == (1, 1, 4)
