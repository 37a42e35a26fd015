"""The CUDA path against the oracle AT BASELINE.json's named shapes (not at reduced sizes):

  configs[1]  paper shape T=240, N=10,000, 46 characteristics, 178 macro series: one full fwd+bwd of
              phase 'conditional' against oracle/torch_port.step (the bit-faithful restatement of the
              reference): losses 1e-5, weights 1e-4, Sharpe, portfolio returns, all twelve gradients;
  configs[2]  full-sample inference shape T=600: get_weights(normalized=True) against the port;
  plus the `weighted_loss=False` configuration (src/model.py:368-369) and a 200-step Adam trajectory
  of the tensor-core path against the port (src/train.py:77-94).

The port needs ~8 GB of host memory and a few seconds per step at the paper shape.
"""
import numpy as np
import pytest
import torch

from _dlap_testlib import rel_err
from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
from oracle import closed_form as cf
from oracle import torch_port as tp
from test_parity_gpu import LOSS_RTOL, WEIGHT_TOL, _np, dev, make_panel

pytestmark = pytest.mark.gpu

PAPER = dict(T=240, N=10_000, M=178, F=46)
TRAJ_TOL = {"fp32": 5e-2, "bf16x3": 2e-1}     # provisional: to be set to ~2.5x the measured values


@pytest.fixture(scope="module")
def paper_case():
    """Paper-shape panel, seed-42 default initialisation, and the port's step on it (computed once)."""
    macro, x, R, mask = make_panel(PAPER["T"], PAPER["N"], PAPER["M"], seed=1234)
    params = tp.init_params(PAPER["M"], PAPER["F"], seed=42)
    out, grads = tp.step(params, macro, x, R, mask, phase="conditional")
    return dict(panel=(macro, x, R, mask), params=params, out=out, grads=grads)


def _grad_errs(got, ref):
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    return (np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30),
            np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-30))


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_paper_shape_step_vs_port(paper_case, precision):
    d = dev()
    macro, x, R, mask = paper_case["panel"]
    torch.manual_seed(42)
    model = AssetPricingGAN({"macro_feature_dim": PAPER["M"], "individual_feature_dim": PAPER["F"],
                             "dropout": 0.0, "b200_precision": precision}).to(d).eval()
    for k, v in model.state_dict().items():          # same seed, same construction order -> same parameters
        assert torch.equal(v.cpu(), paper_case["params"][k]), k
    out = model(macro.to(d), x.to(d), R.to(d), mask.to(d), phase="conditional")
    out["loss"].backward()
    torch.cuda.synchronize()
    ref = paper_case["out"]
    for key in ("loss", "loss_unconditional", "loss_conditional"):
        a, b = float(out[key]), float(ref[key])
        assert abs(a - b) <= LOSS_RTOL * abs(b), (key, a, b)
    w = _np(out["weights"])
    assert np.all(w[~_np(mask)] == 0.0)
    assert rel_err(w, _np(ref["weights"])) < WEIGHT_TOL
    assert rel_err(_np(out["portfolio_returns"]), _np(ref["portfolio_returns"])) < WEIGHT_TOL
    assert abs(float(out["sharpe"]) - float(ref["sharpe"])) <= 1e-4 * abs(float(ref["sharpe"])) + 1e-7
    hn, cn = out["hidden"]
    assert np.abs(_np(hn) - _np(ref["hidden"][0])).max() < 2e-6 and np.abs(_np(cn) - _np(ref["hidden"][1])).max() < 2e-6
    # gradients.  The port itself is fp32 (sums over 2.4 M stock-months in torch's CPU kernels), so it is only
    # good to ~1e-4 of a tensor's norm here; the tensor-core path adds its ReLU-kink noise (test_tc_parity_gpu.py)
    tol_norm = 1e-3 if precision == "fp32" else 1e-2
    checked = 0
    for k, p in model.named_parameters():
        r = paper_case["grads"][k]
        if k == "sdf_net.output_proj.bias":
            assert float(p.grad.abs().max()) == 0.0            # identically zero (demeaning)
            continue
        if k.startswith("moment_net"):
            assert p.grad is not None                           # phase 3 fills both groups
        emax, enorm = _grad_errs(_np(p.grad), _np(r))
        print(f"[{precision}] paper shape, grad {k}: max-rel {emax:.1e}, norm-rel {enorm:.1e}")
        assert enorm < tol_norm and emax < 4 * tol_norm, (k, emax, enorm)
        checked += 1
    assert checked == 11


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_full_sample_normalized_weights_vs_port(precision):
    """configs[2]: ~600 months (train+valid+test in one pass), inference only."""
    d = dev()
    macro, x, R, mask = make_panel(600, 10_000, 178, seed=7)
    params = tp.init_params(178, 46, seed=42)
    with torch.no_grad():
        rw, rhid = tp.normalized_weights(params, macro, x, mask)
        rpr = (rw * R * mask.float()).sum(dim=1)
    torch.manual_seed(42)
    model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46, "dropout": 0.0,
                             "b200_precision": precision}).to(d).eval()
    with torch.no_grad():
        wn, hidden = model.get_weights(macro.to(d), x.to(d), mask.to(d), normalized=True)
        f = model.get_sdf_factor(macro.to(d), x.to(d), R.to(d), mask.to(d))
    assert np.all(_np(wn)[~_np(mask)] == 0.0)
    assert rel_err(_np(wn), _np(rw)) < WEIGHT_TOL
    assert rel_err(_np(f), _np(rpr)) < WEIGHT_TOL
    sa, sb = float(f.mean() / f.std()), float(rpr.mean() / rpr.std())
    assert abs(sa - sb) <= 1e-4 * abs(sb) + 1e-7                                  # out-of-sample Sharpe
    assert np.abs(_np(hidden[0]) - _np(rhid[0])).max() < 2e-6


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
@pytest.mark.parametrize("phase", ["unconditional", "conditional"])
def test_unweighted_loss_matches_oracle(precision, phase):
    """weighted_loss=False: pr_t = sum_i w R m without the N_bar / N_t rescaling (src/model.py:368-369, :412-413)."""
    d = dev()
    macro, x, R, mask = make_panel(9, 700, 8, seed=21)
    mask[4] = False
    x = x * mask[:, :, None]; R = R * mask
    torch.manual_seed(3)
    model = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.0,
                             "weighted_loss": False, "b200_precision": precision}).to(d).eval()
    params = {k: _np(v) for k, v in model.state_dict().items()}
    out = model(macro.to(d), x.to(d), R.to(d), mask.to(d), phase=phase)
    out["loss"].backward()
    torch.cuda.synchronize()
    ref = cf.forward(params, _np(macro), _np(x), _np(R), _np(mask), phase=phase, weighted_loss=False)
    gref = cf.backward(params, _np(macro), _np(x), _np(R), _np(mask), ref, phase=phase, weighted_loss=False)
    chk = cf.forward(params, _np(macro), _np(x), _np(R), _np(mask), phase=phase, weighted_loss=True)
    assert abs(chk["loss"] - ref["loss"]) > 1e-3 * abs(ref["loss"])           # the flag matters on this panel
    # ... and the port (the reference restated) says the same as the closed form
    pout = tp.forward({k: torch.as_tensor(v) for k, v in params.items()}, macro, x, R, mask, phase=phase,
                      weighted_loss=False)
    assert abs(float(pout["loss"]) - ref["loss"]) <= 2e-5 * abs(ref["loss"])
    for key in ("loss", "loss_unconditional", "loss_conditional"):
        a, b = float(out[key]), float(ref[key])
        assert abs(a - b) <= LOSS_RTOL * abs(b) + 1e-30, (key, a, b)
    assert rel_err(_np(out["weights"]), ref["weights"]) < WEIGHT_TOL
    assert rel_err(_np(out["portfolio_returns"]), ref["portfolio_returns"]) < WEIGHT_TOL
    tol = 2e-4 if precision == "fp32" else 1e-2
    for k, p in model.named_parameters():
        if p.grad is None or k == "sdf_net.output_proj.bias":
            continue
        emax, enorm = _grad_errs(_np(p.grad), gref[k].reshape(p.shape))
        assert enorm < tol, (k, emax, enorm)


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_adam_trajectory_200_steps(precision):
    """200 Adam steps (lr 1e-3, clip 1.0, src/train.py:77-94) of the phase-3 loss against the port on the CPU, from
    the same initialisation: the loss curves must stay together, and so must the portfolio (normalised weights) at
    the end.  Two fp32 implementations already differ here: their sums run in different orders, so some ReLU inputs
    within an ulp of zero fall on different sides, and Adam (steps of +-lr whatever the gradient's size) lets such
    differences grow.  The exact-fp32 kernels show that floor; the tensor-core kernels (pre-activations good to 2^-18)
    sit a constant factor above it.  The measured deviations are printed (`pytest -s`)."""
    d = dev()
    macro, x, R, mask = make_panel(60, 2000, 8, seed=31)
    torch.manual_seed(12)
    model = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.0,
                             "b200_precision": precision}).to(d)
    ref = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in model.state_dict().items()}
    opt = torch.optim.Adam(model.sdf_net.parameters(), lr=1e-3)
    ropt = torch.optim.Adam([ref[k] for k in tp.SDF_KEYS], lr=1e-3)
    args = (macro.to(d), x.to(d), R.to(d), mask.to(d))
    worst, at = 0.0, -1
    first = last = None
    for step in range(200):
        opt.zero_grad()
        out = model(*args, phase="conditional")
        out["loss"].backward()
        torch.nn.utils.clip_grad_norm_(model.sdf_net.parameters(), 1.0)
        opt.step()
        ropt.zero_grad()
        rout = tp.forward(ref, macro, x, R, mask, phase="conditional")
        rout["loss"].backward()
        torch.nn.utils.clip_grad_norm_([ref[k] for k in tp.SDF_KEYS], 1.0)
        ropt.step()
        a, b = float(out["loss"]), float(rout["loss"])
        first = b if first is None else first
        last = b
        dev_ = abs(a - b) / max(abs(b), 1e-3 * first)     # relative; below 0.1 % of the initial loss: to that floor
        if dev_ > worst:
            worst, at = dev_, step
    with torch.no_grad():
        model.eval()
        wn, _ = model.get_weights(args[0], args[1], args[3], normalized=True)
        rwn, _ = tp.normalized_weights({k: v.detach() for k, v in ref.items()}, macro, x, mask)
    werr = rel_err(_np(wn), _np(rwn))
    print(f"[{precision}] 200 Adam steps: loss {first:.3e} -> {last:.3e}; worst relative loss deviation {worst:.2e} "
          f"(step {at}); normalised weights after the run {werr:.2e}")
    assert last < 0.9 * first                      # the run actually trained
    assert worst < TRAJ_TOL[precision], worst
    assert werr < 10 * TRAJ_TOL[precision]
