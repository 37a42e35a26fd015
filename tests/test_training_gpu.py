"""clip_grad_norm_ + Adam as one launch, and the whole epoch of src/train.py:45-103 as one CUDA
graph -- against torch.optim.Adam + torch.nn.utils.clip_grad_norm_ themselves."""
import numpy as np
import pytest
import torch

from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, FusedClipAdam, train_epoch
from test_parity_gpu import dev, make_panel

pytestmark = pytest.mark.gpu

SHAPES = [(16, 178), (16, 4), (16,), (16,), (64, 50), (64,), (64, 64), (64,), (1, 64), (1,), (8, 224), (8,)]


@pytest.mark.parametrize("max_norm", [0.0, 1.0])
def test_fused_clip_adam_matches_torch_over_100_steps(max_norm):
    d = dev()
    g = torch.Generator(device=d).manual_seed(5)
    ours = [torch.nn.Parameter(torch.randn(*s, generator=g, device=d)) for s in SHAPES]
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ours]
    opt = FusedClipAdam(ours, lr=1e-3, max_norm=max_norm)
    ropt = torch.optim.Adam(ref, lr=1e-3)
    for step in range(100):
        scale = 10.0 ** float(torch.randint(-4, 1, (1,)))          # gradient norms on both sides of the clip
        for p, r in zip(ours, ref):
            gr = scale * torch.randn(p.shape, generator=g, device=d)
            p.grad = gr.clone(); r.grad = gr.clone()
        if step % 7 == 3:                                            # a parameter without gradient is skipped
            ours[3].grad = None; ref[3].grad = None
        if max_norm > 0:
            rn = torch.nn.utils.clip_grad_norm_(ref, max_norm)
        ropt.step()
        versions = [p._version for p in ours]
        opt.step()
        assert all(p._version > v for p, v in zip(ours, versions) if p.grad is not None)
        if max_norm > 0:
            assert abs(float(opt.grad_norm) - float(rn)) <= 2e-6 * float(rn)
            for p, r in zip(ours, ref):                              # clip_grad_norm_ scales the gradients in place
                if p.grad is not None:
                    assert torch.allclose(p.grad, r.grad, rtol=2e-6, atol=0)
    for p, r in zip(ours, ref):
        err = float((p.detach() - r.detach()).abs().max() / r.detach().abs().max())
        assert err < 1e-6, err
    sd = opt.state_dict()
    rsd = ropt.state_dict()
    assert float(sd["state"][0]["step"]) == float(rsd["state"][0]["step"]) == 100.0
    assert float(sd["state"][3]["step"]) == float(rsd["state"][3]["step"]) < 100.0      # per-parameter step counts
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max())
    e_m = max(rel(sd["state"][i]["exp_avg"], rsd["state"][i]["exp_avg"]) for i in range(len(SHAPES)))
    e_v = max(rel(sd["state"][i]["exp_avg_sq"], rsd["state"][i]["exp_avg_sq"]) for i in range(len(SHAPES)))
    print(f"[max_norm={max_norm}] after 100 steps vs torch.optim.Adam: exp_avg {e_m:.1e}, exp_avg_sq {e_v:.1e}")
    assert e_m < 1e-5 and e_v < 1e-4


@pytest.mark.parametrize("precision,phase,dropout", [("bf16x3", "conditional", 0.0), ("fp32", "unconditional", 0.0),
                                                     ("bf16x3", "moment", 0.0)])
def test_graphed_epoch_matches_the_eager_torch_loop(precision, phase, dropout):
    """zero_grad + forward + backward + clip + Adam replayed from one CUDA graph against the same model stepped
    eagerly with torch.optim.Adam and clip_grad_norm_ (src/train.py:77-94): 20 epochs."""
    d = dev()
    macro, x, R, mask = [t.to(d) for t in make_panel(20, 600, 8, seed=3)]
    cfg = {"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": dropout, "b200_precision": precision}
    torch.manual_seed(9)
    a = AssetPricingGAN(cfg).to(d)
    b = AssetPricingGAN(cfg).to(d)
    b.load_state_dict(a.state_dict())
    scope = "moment" if phase == "moment" else "sdf"
    for m in (a, b):
        for p in m.sdf_net.parameters():
            p.requires_grad = phase != "moment"
    pa = list(a.moment_net.parameters() if scope == "moment" else a.sdf_net.parameters())
    pb = list(b.moment_net.parameters() if scope == "moment" else b.sdf_net.parameters())
    ropt = torch.optim.Adam(pa, lr=1e-3)
    opt = FusedClipAdam(pb, lr=1e-3)
    before = [p.detach().clone() for p in pb]
    gs = b.train().graphed_step(macro, x, R, mask, phase=phase, optimizer=opt, max_norm=1.0)
    assert all(torch.equal(p, q) for p, q in zip(pb, before))          # warm-up and capture moved nothing
    # per-cell forward (+ the LSTM of its own for fp32), reductions, per-cell backward, month-level backward, clip + Adam;
    # phase 2 freezes the SDF net: no BPTT kernel for the month-summed gradients to ride in, so they are one launch more
    assert gs.launches == (5 if precision == "bf16x3" else 6) + (1 if phase == "moment" else 0)
    a.train()
    for epoch in range(20):
        ropt.zero_grad()
        out = a(macro, x, R, mask, phase=phase)
        out["loss"].backward()
        rn = torch.nn.utils.clip_grad_norm_(pa, 1.0)
        ropt.step()
        got = gs.step()
        torch.cuda.synchronize()
        # two copies of the model stepped by two Adam implementations: equal to rounding, which Adam's +-lr steps
        # amplify a little every epoch
        assert abs(float(got["loss"].detach()) - float(out["loss"].detach())) <= 2e-3 * abs(float(out["loss"].detach())), epoch
        assert abs(float(opt.grad_norm) - float(rn)) <= 2e-3 * float(rn)
        rep = gs.report.cpu().numpy()
        assert rep[0] == float(got["loss"]) and rep[3] == float(opt.grad_norm)
        assert rep[4] == 0.0                                      # residual loss: off in this configuration
        assert np.array_equal(rep[5:], got["portfolio_returns"].detach().cpu().numpy())
    for (k, p), q in zip(a.named_parameters(), b.parameters()):
        if k == "sdf_net.output_proj.bias":
            continue
        # (norm-wise: an element whose gradient hovers around zero takes +-lr steps of either sign in the two runs)
        assert float((p - q).norm()) <= 5e-3 * float(p.norm()), k


def test_train_epoch_signature_and_report():
    """train_epoch(model, optimizer, data, device, phase, grad_clip, scope) -> the reference's dict, from host tensors
    that are uploaded once."""
    d = dev()
    macro, x, R, mask = make_panel(12, 300, 8, seed=4)
    data = {"macro_features": macro, "individual_features": x, "returns": R, "mask": mask}
    torch.manual_seed(2)
    model = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.0}).to(d)
    ref = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46, "dropout": 0.0}).to(d)
    ref.load_state_dict(model.state_dict())
    opt = FusedClipAdam(model.sdf_net.parameters(), lr=1e-3)
    ropt = torch.optim.Adam(ref.sdf_net.parameters(), lr=1e-3)
    for epoch in range(5):
        m = train_epoch(model, opt, data, d, phase="unconditional", scope="sdf")
        assert sorted(m) == ["grad_norm", "loss", "loss_cond", "loss_residual", "loss_unc", "sharpe"]
        ref.train(); ropt.zero_grad()
        out = ref(macro.to(d), x.to(d), R.to(d), mask.to(d), phase="unconditional")
        out["loss"].backward()
        rn = torch.nn.utils.clip_grad_norm_(ref.sdf_net.parameters(), 1.0)
        ropt.step()
        pr = out["portfolio_returns"].detach()
        assert abs(m["loss"] - float(out["loss"])) <= 2e-5 * abs(float(out["loss"]))
        assert abs(m["sharpe"] - float(pr.mean() / pr.std())) <= 1e-4 * abs(float(pr.mean() / pr.std())) + 1e-6
        assert abs(m["grad_norm"] - float(rn)) <= 2e-5 * float(rn) and m["loss_cond"] == 0.0
    assert len(model._epoch_graphs) == 1                              # one capture, five replays
    with pytest.raises(NotImplementedError):
        train_epoch(model, opt, data, d, phase="conditional", scope="all")
    with pytest.raises(TypeError):
        train_epoch(model, ropt, data, d)
