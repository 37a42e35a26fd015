"""Device-side loader against the reference's loader, bit for bit.

Goldens: tests/golden/loader.npz, written by the reference's own AssetPricingDataset / create_data_loaders /
create_small_sample (tests/golden/make_golden.py `loader`).  Where the unmodified reference is staged
(baseline/_ref, by build()), its loader is also run here on the shipped files and compared directly."""
import os
import sys

import numpy as np
import pytest
import torch

from _dlap_testlib import load_golden

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bytes(t):
    return (t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)).tobytes()


@pytest.fixture(scope="module")
def D():
    from deeplearninginassetpricing_paperreplication_b200 import data
    return data


@pytest.fixture(scope="module")
def g():
    return load_golden("loader")


@pytest.mark.parametrize("source", ["host", "device"])
@pytest.mark.parametrize("split", ["train", "valid", "f64"])
def test_ingest_panel_matches_reference_loader(D, g, split, source):
    """src/data_loader.py:45-65: sentinels, NaNs, values on the threshold, an empty month, -0.0, fp64 files; the raw
    array read out of page-locked host memory by the kernel, or already on the device."""
    raw = g[f"{split}/raw"]
    src = torch.from_numpy(raw).cuda() if source == "device" else raw
    R, x, mask = D.ingest_panel(src, "cuda")
    assert mask.dtype == torch.bool and R.dtype == x.dtype == torch.float32
    assert np.array_equal(mask.cpu().numpy(), g[f"{split}/mask"])
    assert _bytes(R) == g[f"{split}/returns"].tobytes()
    assert _bytes(x) == g[f"{split}/features"].tobytes()


def test_zscore_matches_numpy_order(D, g):
    """src/data_loader.py:72-86: statistics of the training file (with macro_idx: numpy sums pairwise; without: row
    after row), the validation file normalised with them, a constant column (std = 1e-8), 300 months."""
    idx = g["macro_idx"]
    z, mean, std = D.zscore_macro(g["train/macro_raw"][:, idx], device="cuda", pairwise=True)
    assert _bytes(mean) == g["train/mean"].tobytes() and _bytes(std) == g["train/std"].tobytes()
    assert _bytes(z) == g["train/macro"].tobytes()
    zv, mv, sv = D.zscore_macro(g["valid/macro_raw"][:, idx], mean, std, device="cuda")
    assert _bytes(zv) == g["valid/macro"].tobytes() and _bytes(mv) == g["train/mean"].tobytes()
    z, mean, std = D.zscore_macro(g["train/macro_raw"], device="cuda")
    assert _bytes(z) == g["train/macro_all"].tobytes() and _bytes(std) == g["train/std_all"].tobytes()
    z, mean, std = D.zscore_macro(g["long/macro_raw"][:, g["long/idx"]], device="cuda", pairwise=True)
    assert _bytes(z) == g["long/macro"].tobytes() and _bytes(mean) == g["long/mean"].tobytes()
    assert _bytes(std) == g["long/std"].tobytes()
    z, mean, std = D.zscore_macro(g["long/macro_raw"], device="cuda")
    assert _bytes(z) == g["long/macro_all"].tobytes() and _bytes(mean) == g["long/mean_all"].tobytes()


def test_dataset_classes_from_files(D, g, tmp_path):
    """The mirror of AssetPricingDataset / create_data_loaders / create_small_sample on .npz files."""
    paths = {}
    for split in ("train", "valid"):
        paths[split] = (str(tmp_path / f"char_{split}.npz"), str(tmp_path / f"macro_{split}.npz"))
        np.savez(paths[split][0], data=g[f"{split}/raw"])
        np.savez(paths[split][1], data=g[f"{split}/macro_raw"])
    idx = [int(i) for i in g["macro_idx"]]
    tr, va, te = D.create_data_loaders(*paths["train"], *paths["valid"], macro_idx=idx)
    assert te is None
    for split, ds in (("train", tr), ("valid", va)):
        b = ds.get_full_batch()
        assert all(v.is_cuda for v in b.values())
        assert np.array_equal(b["mask"].cpu().numpy(), g[f"{split}/mask"])
        assert _bytes(b["returns"]) == g[f"{split}/returns"].tobytes()
        assert _bytes(b["individual_features"]) == g[f"{split}/features"].tobytes()
        assert _bytes(b["macro_features"]) == g[f"{split}/macro"].tobytes()
        assert np.array_equal(ds.get_date_count_list(), g[f"{split}/date_count"])
        assert (ds.T, ds.N) == g[f"{split}/mask"].shape and len(ds) == ds.T
        assert ds.individual_feature_dim == 46 and ds.macro_feature_dim == len(idx)
        item = ds[3]
        assert _bytes(item["returns"]) == g[f"{split}/returns"][3].tobytes()
    assert _bytes(va.mean_macro) == g["train/mean"].tobytes()
    small = D.create_small_sample(tr, n_periods=12, n_stocks=40)
    for k, v in small.items():
        assert _bytes(v) == g["small/" + k].tobytes(), k
    bare = D.AssetPricingDataset(paths["valid"][0])                 # no macro file
    assert bare.macro_features is None and bare.macro_feature_dim == 0 and "macro_features" not in bare.get_full_batch()
    raw_macro = D.AssetPricingDataset(*paths["valid"], normalize_macro=False)
    assert _bytes(raw_macro.macro_features) == g["valid/macro_raw"].tobytes() and raw_macro.mean_macro is None


@pytest.mark.parametrize("T,N,F", [(240, 10000, 46), (7, 1001, 46), (3, 129, 7), (5, 77, 64), (2, 300, 1), (1, 1, 46)])
def test_ingest_full_size_against_the_same_logic_in_torch(D, T, N, F):
    """BASELINE.json's paper shape and ragged sizes (chunk tails, F+1 even / odd, one cell): the kernel against the
    loader's expressions evaluated by torch on the same device.  Comparisons and copies only: bit-exact."""
    gen = torch.Generator(device="cuda").manual_seed(T * 1000 + N + F)
    raw = torch.randn(T, N, F + 1, device="cuda", generator=gen)
    miss = torch.rand(T, N, device="cuda", generator=gen)
    raw[miss < 0.35] = -99.99
    raw[..., 0][(miss > 0.35) & (miss < 0.40)] = -99.99
    one = (miss > 0.40) & (miss < 0.45)
    col = torch.randint(1, F + 1, (T, N), device="cuda", generator=gen)
    raw[one] = raw[one].scatter(1, col[one][:, None], -99.99)
    raw[0, 0, 0] = float("nan") if T * N > 1 else raw[0, 0, 0]
    thr = D.VALID_ABOVE
    want_mask = (raw[..., 0] > thr) & ~torch.isnan(raw[..., 0]) & (raw[..., 1:] > thr).all(dim=2)
    want_R = torch.where(want_mask, raw[..., 0], torch.zeros((), device="cuda"))
    want_x = torch.where(want_mask[..., None], raw[..., 1:], torch.zeros((), device="cuda"))
    R, x, mask = D.ingest_panel(raw, "cuda")
    assert torch.equal(mask, want_mask)
    assert torch.equal(R.view(torch.int32), want_R.view(torch.int32))
    assert torch.equal(x.view(torch.int32), want_x.contiguous().view(torch.int32))
    frac = mask.float().mean().item()
    assert T * N < 1000 or 0.4 < frac < 0.7


def test_reference_loader_on_the_shipped_files(D):
    """The unmodified reference's loader (staged under baseline/_ref) on its shipped files, next to ours."""
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    data_dir = os.path.join(ref_root, "data", "synthetic_data")
    if not os.path.isdir(os.path.join(ref_root, "src")) or not os.path.isdir(data_dir):
        pytest.skip("baseline/_ref is not staged (build() stages it where /root/reference exists)")
    sys.path.insert(0, ref_root)
    try:
        from src.data_loader import create_data_loaders as ref_loaders
    finally:
        sys.path.remove(ref_root)
    p = lambda kind, split: os.path.join(data_dir, kind, f"{'Char' if kind == 'char' else 'macro'}_{split}.npz")
    args = (p("char", "train"), p("macro", "train"), p("char", "valid"), p("macro", "valid"), p("char", "test"),
            p("macro", "test"))
    for ours, ref in zip(D.create_data_loaders(*args), ref_loaders(*args)):
        assert np.array_equal(ours.mask.cpu().numpy(), ref.mask)
        assert _bytes(ours.returns) == ref.returns.tobytes()
        assert _bytes(ours.individual_features) == ref.individual_features.tobytes()
        assert _bytes(ours.macro_features) == ref.macro_features.tobytes()
        assert _bytes(ours.mean_macro) == ref.mean_macro.tobytes() and _bytes(ours.std_macro) == ref.std_macro.tobytes()


def test_step_on_a_loaded_dataset_uses_the_index_built_at_load_time(D, g, tmp_path):
    """get_full_batch() hands CUDA tensors to the unchanged training loop; the valid-cell index made by the loader
    is the one the first step finds."""
    from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN, functional as Fn
    cp, mp = str(tmp_path / "c.npz"), str(tmp_path / "m.npz")
    np.savez(cp, data=g["train/raw"]); np.savez(mp, data=g["train/macro_raw"])
    Fn.invalidate_panel_index()
    ds = D.AssetPricingDataset(cp, mp)
    b = ds.get_full_batch()
    assert any(hit[0]() is b["mask"] for hit in Fn._index_cache)
    torch.manual_seed(3)
    model = AssetPricingGAN({"macro_feature_dim": ds.macro_feature_dim, "individual_feature_dim": 46}).cuda().eval()
    dev = torch.device("cuda")
    out = model(b["macro_features"].to(dev), b["individual_features"].to(dev), b["returns"].to(dev), b["mask"].to(dev),
                phase="conditional")
    n_idx = len(Fn._index_cache)
    clone = model(b["macro_features"].clone(), b["individual_features"].clone(), b["returns"].clone(),
                  b["mask"].clone(), phase="conditional")
    assert len(Fn._index_cache) == n_idx + 1 or len(Fn._index_cache) == 4      # the clone's mask needed a new index
    assert float(out["loss"]) == float(clone["loss"]) and torch.equal(out["weights"], clone["weights"])
    assert torch.isfinite(out["loss"])
