"""X images (dlap_panel_image): the characteristics in the tensor-core kernels' operand format, built once per panel.

The kernels must produce the same bits whether they fetch a tile's image with one bulk copy or gather and convert
its rows themselves; the image itself is checked against numpy (layout, hi + lo = x, ones column, repeated last
row); and the cache must follow the panel (second use builds, new values rebuild, a captured graph is kept current)."""
import numpy as np
import pytest
import torch

from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, functional as Fn
from test_parity_gpu import _np, dev, make_panel

pytestmark = pytest.mark.gpu

X_LBO, X_IMG = 128 * 16 + 16, 6 * (128 * 16 + 16)


def _bf16_to_f32(u16):
    return (u16.astype(np.uint32) << 16).view(np.float32)


def _model(M=178, F=46, seed=0, **cfg):
    torch.manual_seed(seed)
    return AssetPricingGAN(dict({"macro_feature_dim": M, "individual_feature_dim": F, "b200_precision": "bf16x3"},
                                **cfg)).to(dev())


def _step(model, panel, phase="conditional"):
    model.zero_grad(set_to_none=True)
    out = model(*panel, phase=phase)
    out["loss"].backward()
    torch.cuda.synchronize()
    res = {k: out[k].detach().clone() for k in ("loss", "loss_unconditional", "loss_conditional", "weights",
                                                 "portfolio_returns", "sharpe")}
    res.update({"grad/" + k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None})
    del out
    return res


def _same(a, b):
    assert a.keys() == b.keys()
    for k in a:
        assert torch.equal(a[k], b[k]), k


@pytest.mark.parametrize("T,N,F", [(9, 1000, 46), (5, 333, 46), (6, 700, 7), (4, 257, 45)])
def test_image_matches_numpy(T, N, F):
    """Every tile of the index: bf16 hi + lo reproduces x to 2^-16, hi is the rounded value, k = 46 holds 1.0, k = 47
    and the columns past F hold 0, rows past the end of a tile repeat its last row."""
    d = dev()
    _, x, _, mask = make_panel(T, N, M=0, F=F, seed=T + N)
    mask[1] = False                                   # an empty month
    x, mask = x.to(d), mask.to(d)
    ws = Fn.PlanOnly(T, N, F, d)
    index = Fn.panel_index(ws, mask)
    Fn.forget_panel_image(x)
    image = Fn.panel_image(ws, x, index, force=True)
    torch.cuda.synchronize()
    idx = index.cpu().numpy()
    ntiles = int(idx[0])
    assert image.numel() == max(ntiles, 1) * 2 * X_IMG
    ts = idx[4:4 + T + 1]
    cnt = idx[4 + T + 1:4 + 2 * T + 1]
    voff = (4 + 2 * T + 1 + 3) & ~3
    vidx = idx[voff:voff + T * N].reshape(T, N)
    assert np.array_equal(cnt, mask.sum(dim=1).cpu().numpy()) and ts[T] == ntiles
    img = image.cpu().numpy().view(np.uint16).reshape(ntiles, 2, 6, X_LBO // 2)
    xs = x.cpu().numpy()
    for t in range(T):
        for k in range(ts[t], ts[t + 1]):
            j0 = (k - ts[t]) * 128
            rows = np.minimum(j0 + np.arange(128), cnt[t] - 1)
            want = np.zeros((128, 48), np.float32)
            want[:, :F] = xs[t, vidx[t, rows]]
            want[:, 46] = 1.0
            hi = _bf16_to_f32(img[k, 0, :, :128 * 8].reshape(6, 128, 8)).transpose(1, 0, 2).reshape(128, 48)
            lo = _bf16_to_f32(img[k, 1, :, :128 * 8].reshape(6, 128, 8)).transpose(1, 0, 2).reshape(128, 48)
            assert np.array_equal(hi[:, 46], np.ones(128, np.float32)) and not lo[:, 46].any()
            assert not hi[:, 47].any() and not hi[:, F:46].any() and not lo[:, F:46].any()
            assert np.abs(hi + lo - want).max() <= 2.0 ** -16 * np.abs(want).max()
            assert np.abs(hi - want).max() <= 2.0 ** -8 * np.abs(want).max()


@pytest.mark.parametrize("phase", ["unconditional", "moment", "conditional"])
@pytest.mark.parametrize("T,N", [(12, 1500), (7, 129)])
def test_step_is_bit_identical_with_and_without_the_image(T, N, phase, monkeypatch):
    d = dev()
    panel = [t.to(d) for t in make_panel(T, N, seed=3)]
    panel[3][2] = False
    panel[1][2] = 0
    model = _model().eval()
    for p in model.sdf_net.parameters():
        p.requires_grad = phase != "moment"
    monkeypatch.setattr(Fn, "use_x_image", False)
    gathered = _step(model, panel, phase)
    assert Fn.cached_panel_image(panel[1]) is None
    monkeypatch.setattr(Fn, "use_x_image", True)
    monkeypatch.setattr(Fn, "force_x_image", True)
    imaged = _step(model, panel, phase)
    assert Fn.cached_panel_image(panel[1]) is not None
    _same(gathered, imaged)


def test_dropout_is_bit_identical_with_and_without_the_image(monkeypatch):
    d = dev()
    panel = [t.to(d) for t in make_panel(10, 900, seed=5)]
    model = _model(dropout=0.1).train()
    seed = torch.tensor([1234567], dtype=torch.int64, device=d)
    model._seed_dev = seed
    try:
        monkeypatch.setattr(Fn, "use_x_image", False)
        gathered = _step(model, panel)
        monkeypatch.setattr(Fn, "use_x_image", True)
        monkeypatch.setattr(Fn, "force_x_image", True)
        imaged = _step(model, panel)
    finally:
        model._seed_dev = None
    _same(gathered, imaged)


def test_image_follows_the_panel():
    """First sight of a panel: no image (a loop that uploads a fresh panel per step never pays for one); second
    step on the same tensor: built; new values (version counter): dropped, then rebuilt; results track the values."""
    d = dev()
    macro, x, R, mask = [t.to(d) for t in make_panel(8, 600, seed=7)]
    model = _model().eval()
    Fn.invalidate_panel_index()
    first = _step(model, (macro, x, R, mask))
    assert Fn.cached_panel_image(x) is None
    second = _step(model, (macro, x, R, mask))
    assert Fn.cached_panel_image(x) is not None
    _same(first, second)
    x.mul_(0.5)
    half = _step(model, (macro, x, R, mask))
    assert Fn.cached_panel_image(x) is None and not torch.equal(half["loss"], first["loss"])
    again = _step(model, (macro, x, R, mask))
    assert Fn.cached_panel_image(x) is not None
    _same(half, again)
    fresh = _step(model, (macro, x.clone(), R, mask))
    _same(half, fresh)
    Fn.forget_panel_image(x)
    assert Fn.cached_panel_image(x) is None


def test_graphed_step_rebuilds_the_image_when_the_panel_gets_new_values():
    d = dev()
    macro, x, R, mask = [t.to(d) for t in make_panel(8, 700, seed=11)]
    model = _model().eval()
    gs = model.graphed_step(macro, x, R, mask, phase="conditional")
    assert gs._ximg is not None
    out = gs.step()
    torch.cuda.synchronize()
    eager = _step(_model().eval(), (macro, x.clone(), R, mask))
    assert torch.equal(out["loss"].detach(), eager["loss"])
    x.copy_(x * 0.25)
    out = gs.step()
    torch.cuda.synchronize()
    loss_new = out["loss"].detach().clone()
    grad_new = model.sdf_net.fc_layers[0].weight.grad.detach().clone()
    eager = _step(_model().eval(), (macro, x.clone(), R, mask))
    assert torch.equal(loss_new, eager["loss"])
    assert torch.equal(grad_new, eager["grad/sdf_net.fc_layers.0.weight"])


def test_paper_shape_image_and_gather_agree(monkeypatch):
    """BASELINE.json configs[1]: T=240, N=10,000 -- 7,500 tiles, every CTA's range, the fused LSTM beside them."""
    import bench
    d = dev()
    panel = [t.to(d) for t in bench.synthetic_panel(240, 10000, 46, 178, seed=99)]
    model = _model().eval()
    monkeypatch.setattr(Fn, "use_x_image", False)
    gathered = _step(model, panel)
    monkeypatch.setattr(Fn, "use_x_image", True)
    monkeypatch.setattr(Fn, "force_x_image", True)
    imaged = _step(model, panel)
    _same(gathered, imaged)
    index, image = Fn.cached_panel_image(panel[1])
    assert image.numel() == int(index[0].item()) * 2 * X_IMG
