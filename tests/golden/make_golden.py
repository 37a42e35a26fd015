#!/usr/bin/env python
"""Regenerate the golden fixtures in this directory FROM THE REFERENCE ITSELF.

Run in the build container only (it imports ``/root/reference``, which does not exist on the
GPU box):

    python tests/golden/make_golden.py

For every case it builds the unmodified ``AssetPricingGAN`` (``src/model.py:298``) under
``torch.manual_seed``, puts it in ``eval()`` (dropout disabled, as every parity comparison
requires), runs ``forward`` + ``loss.backward()`` for the three phases exactly as
``train_epoch`` does (``src/train.py:77-84``; phase 2 with the SDF net frozen as at
``src/train.py:309-312``), plus ``get_weights(normalized=True)`` (``src/model.py:565``), and
stores inputs, parameters and every output as float32 ``.npz``.

Cases
  rand_small    T=12, N=37 (odd), 178 macros: one empty month, one never-valid stock.
  shipped_sub   the shipped seed-42 panel through ``AssetPricingDataset`` (all 120 months,
                96 of its 500 stocks, 8 macros) -- exercises the loader's mask contract.
  shipped_valid the shipped validation split (30 months, 160 stocks incl. never-valid ones),
                macro z-scored with TRAIN statistics (``src/train.py:485-497``), non-zero
                initial LSTM state.
  mask_raw      raw returns/features with -99.99 sentinels and NaNs -> mask, zero-filled
                arrays (``src/data_loader.py:50-65``).  Bit-exact contract.
  loader        the reference's loader (``src/data_loader.py:42-99, 157-237``) on a raw panel with sentinels, NaNs and
                a wide-range macro series written to temporary .npz files: mask, zero-filled arrays, macro mean / std /
                z-scores of a training file and of a validation file normalised with the training statistics,
                ``get_date_count_list`` and ``create_small_sample`` -- `python make_golden.py loader` writes only this.
                Bit-exact contract for the device-side loader.
  anchors       scalars of the full shipped training panel.
  v_residual    residual_loss_factor = 0.25 (src/model.py:435-483): the regulariser on the weights, a month with one
                listed stock -- `python make_golden.py residual` writes only this.
  v_*           config variants: the notebooks' [64,32] / use_rnn=False / hidden_dim_moment=[32],
                normalize_w=False, narrow hidden layers with weighted_loss=False -- `python make_golden.py variants`
                writes only these.
"""
import json
import os
import sys

import numpy as np
import torch

REF = os.environ.get("DLAP_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
from src.model import AssetPricingGAN  # noqa: E402
from src.data_loader import AssetPricingDataset  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
PHASES = ("unconditional", "moment", "conditional")


def run_case(name, macro, x, R, mask, seed, hidden=None, cfg_extra=None):
    cfg = {"macro_feature_dim": int(macro.shape[1]), "individual_feature_dim": int(x.shape[2])}
    cfg.update(cfg_extra or {})
    torch.manual_seed(seed)
    model = AssetPricingGAN(cfg)
    model.eval()
    blob = {"macro": macro.numpy(), "x": x.numpy(), "R": R.numpy(), "mask": mask.numpy(),
            "seed": np.int64(seed)}
    if cfg_extra:
        blob["cfg_json"] = np.array(json.dumps(cfg_extra))
    if hidden is not None:
        blob["h0"] = hidden[0].numpy(); blob["c0"] = hidden[1].numpy()
    for k, v in model.state_dict().items():
        blob["param/" + k] = v.numpy().copy()
    for phase in PHASES:
        model.zero_grad(set_to_none=True)
        frozen = phase == "moment"
        for prm in model.sdf_net.parameters():
            prm.requires_grad = not frozen
        out = model(macro, x, R, mask, hidden=hidden, phase=phase)
        out["loss"].backward()
        for key in ("weights", "loss", "loss_unconditional", "loss_conditional", "loss_residual", "sharpe",
                    "portfolio_returns"):
            blob[f"{phase}/{key}"] = out[key].detach().numpy().copy()
        if out["hidden"] is not None:
            blob[f"{phase}/hidden_h"] = out["hidden"][0].detach().numpy().copy()
            blob[f"{phase}/hidden_c"] = out["hidden"][1].detach().numpy().copy()
        if phase == "conditional":
            blob["moments"] = out["moments"].detach().contiguous().numpy().copy()
        for k, prm in model.named_parameters():
            if prm.grad is not None:
                blob[f"{phase}/grad/{k}"] = prm.grad.numpy().copy()
    for prm in model.sdf_net.parameters():
        prm.requires_grad = True
    with torch.no_grad():
        wn, _ = model.get_weights(macro, x, mask, hidden=hidden, normalized=True)
        blob["normalized_weights"] = wn.numpy().copy()
        blob["sdf_factor"] = model.get_sdf_factor(macro, x, R, mask, hidden=hidden).numpy().copy()
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **blob)
    print(f"{name}: T={x.shape[0]} N={x.shape[1]} M={macro.shape[1]} "
          f"valid={int(mask.sum())} -> {os.path.getsize(path) / 1024:.0f} KiB")


def rand_small():
    g = torch.Generator().manual_seed(1234)
    T, N, F, M = 12, 37, 46, 178
    mask = torch.rand(T, N, generator=g) < 0.4
    mask[3, :] = False          # empty month  -> clamp(min=1) on N_t
    mask[:, 5] = False          # dead stock   -> clamp(min=1) on T_i, still in the /N mean
    x = torch.randn(T, N, F, generator=g).clamp_(-2, 2) * mask[:, :, None]
    R = 0.1 * torch.randn(T, N, generator=g) * mask
    macro = torch.randn(T, M, generator=g)
    run_case("rand_small", macro, x, R, mask, seed=42)


VARIANTS = {
    "v_notebook": {"hidden_dim": [64, 32], "hidden_dim_moment": [32], "use_rnn": False},   # notebooks/demo*.ipynb
    "v_6432": {"hidden_dim": [64, 32]},
    "v_norm_off": {"normalize_w": False},
    "v_no_rnn": {"use_rnn": False},
    "v_mom_hidden": {"hidden_dim_moment": [32]},
    "v_narrow": {"hidden_dim": [32, 16], "weighted_loss": False},
}


def variants():
    """The configurations of SURVEY.md §8f row 3 (the notebooks' among them) on one small panel."""
    g = torch.Generator().manual_seed(77)
    T, N, F, M = 10, 150, 46, 8
    mask = torch.rand(T, N, generator=g) < 0.45
    mask[6, :] = False
    mask[:, 11] = False
    x = torch.randn(T, N, F, generator=g).clamp_(-2, 2) * mask[:, :, None]
    R = 0.1 * torch.randn(T, N, generator=g) * mask
    macro = torch.randn(T, M, generator=g)
    for i, (name, extra) in enumerate(VARIANTS.items()):
        run_case(name, macro, x, R, mask, seed=100 + i, cfg_extra=extra)


def residual_case():
    """residual_loss_factor > 0 (src/model.py:435-483, :543-548) on the variants' panel, with one month left with a
    single listed stock (skipped by the `valid.sum() < 2` rule) next to the empty one -- `python make_golden.py
    residual` writes only this."""
    g = torch.Generator().manual_seed(77)
    T, N, F, M = 10, 150, 46, 8
    mask = torch.rand(T, N, generator=g) < 0.45
    mask[6, :] = False
    mask[:, 11] = False
    x = torch.randn(T, N, F, generator=g).clamp_(-2, 2)
    R = 0.1 * torch.randn(T, N, generator=g)
    macro = torch.randn(T, M, generator=g)
    mask[3, :] = False
    mask[3, 42] = True
    x = x * mask[:, :, None]
    R = R * mask
    run_case("v_residual", macro, x, R, mask, seed=131, cfg_extra={"residual_loss_factor": 0.25})


def shipped():
    d = os.path.join(REF, "data", "synthetic_data")
    tr = AssetPricingDataset(os.path.join(d, "char", "Char_train.npz"),
                             os.path.join(d, "macro", "macro_train.npz"))
    b = tr.get_full_batch()
    keep = np.arange(0, 480, 5)     # 96 stocks spread over the panel
    run_case("shipped_sub", b["macro_features"], b["individual_features"][:, keep].contiguous(),
             b["returns"][:, keep].contiguous(), b["mask"][:, keep].contiguous(), seed=7)
    mean, std = tr.get_macro_stats()
    va = AssetPricingDataset(os.path.join(d, "char", "Char_valid.npz"),
                             os.path.join(d, "macro", "macro_valid.npz"),
                             mean_macro=mean, std_macro=std)
    vb = va.get_full_batch()
    keep = np.arange(0, 480, 3)
    g = torch.Generator().manual_seed(5)
    hidden = (0.3 * torch.randn(1, 1, 4, generator=g), 0.3 * torch.randn(1, 1, 4, generator=g))
    run_case("shipped_valid", vb["macro_features"],
             vb["individual_features"][:, keep].contiguous(), vb["returns"][:, keep].contiguous(),
             vb["mask"][:, keep].contiguous(), seed=11, hidden=hidden)
    # anchors on the full training panel
    torch.manual_seed(42)
    model = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46}).eval()
    out = model(b["macro_features"], b["individual_features"], b["returns"], b["mask"])
    anchors = {"loss_unconditional": float(out["loss_unconditional"]),
               "loss_conditional": float(out["loss_conditional"]),
               "sharpe": float(out["sharpe"]),
               "portfolio_returns_head": [float(v) for v in out["portfolio_returns"][:3]],
               "abs_weight_sum": float(out["weights"].abs().sum()),
               "mask_sum": int(b["mask"].sum()),
               "torch": torch.__version__}
    with open(os.path.join(HERE, "anchors_shipped_train.json"), "w") as f:
        json.dump(anchors, f, indent=1)
    print("anchors:", anchors)
    # raw mask contract: a slab of the raw file with NaNs injected
    raw = np.load(os.path.join(d, "char", "Char_train.npz"))["data"][:16, :64].astype(np.float32)
    raw = raw.copy()
    raw[2, 3, 0] = np.nan          # NaN return
    raw[4, 7, 9] = np.nan          # NaN feature (fails the '>' test)
    raw[5, 1, 0] = -98.99          # exactly on the threshold (not valid: strict '>')
    raw[6, 2, 4] = -98.98
    tmp = "/tmp/_dlap_mask_raw.npz"
    np.savez(tmp, data=raw)
    ds = AssetPricingDataset(tmp)
    np.savez_compressed(os.path.join(HERE, "mask_raw.npz"), raw=raw, mask=ds.mask,
                        returns=ds.returns, features=ds.individual_features)
    print("mask_raw: valid", int(ds.mask.sum()), "of", ds.mask.size)


def loader():
    """Raw files -> the reference's AssetPricingDataset / create_data_loaders / create_small_sample."""
    from src.data_loader import create_data_loaders, create_small_sample
    rng = np.random.default_rng(20240611)
    files = {}
    for split, (T, N) in {"train": (24, 157), "valid": (7, 157)}.items():
        F = 46
        raw = rng.standard_normal((T, N, F + 1)).astype(np.float32)
        raw[..., 0] *= 0.1
        raw[rng.random((T, N)) < 0.30] = -99.99              # whole stock-month missing
        raw[rng.random((T, N)) < 0.05, 0] = -99.99            # return missing only
        hit = rng.random((T, N)) < 0.05                       # one characteristic missing
        raw[hit, rng.integers(1, F + 1, size=int(hit.sum()))] = -99.99
        raw[1, 2, 0] = np.nan; raw[2, 5, 17] = np.nan         # NaNs fail the comparison
        raw[3, 1, 0] = -98.99; raw[3, 4, 9] = -98.99           # on the threshold: strict '>' says missing
        raw[3, 6, 0] = -98.98; raw[4, 0, 0] = -0.0             # just above; a negative zero survives where valid
        raw[5, :, :] = -99.99                                  # an empty month
        macro = (rng.standard_normal((T, 11)) * rng.uniform(1e-3, 1e3, (1, 11)) + rng.uniform(-50, 50, (1, 11))).astype(np.float32)
        macro[:, 7] = 3.25                                     # a constant column: std = 1e-8
        files[split] = (raw, macro)
        np.savez(f"/tmp/_dlap_loader_char_{split}.npz", data=raw)
        np.savez(f"/tmp/_dlap_loader_macro_{split}.npz", data=macro)
    idx = [0, 2, 3, 7, 10]
    tr, va, _ = create_data_loaders("/tmp/_dlap_loader_char_train.npz", "/tmp/_dlap_loader_macro_train.npz",
                                    "/tmp/_dlap_loader_char_valid.npz", "/tmp/_dlap_loader_macro_valid.npz",
                                    macro_idx=idx)
    blob = {"macro_idx": np.array(idx)}
    for split, ds in (("train", tr), ("valid", va)):
        raw, macro = files[split]
        blob.update({f"{split}/raw": raw, f"{split}/macro_raw": macro, f"{split}/mask": ds.mask,
                     f"{split}/returns": ds.returns, f"{split}/features": ds.individual_features,
                     f"{split}/macro": ds.macro_features, f"{split}/mean": ds.mean_macro, f"{split}/std": ds.std_macro,
                     f"{split}/date_count": ds.get_date_count_list()})
    full = AssetPricingDataset("/tmp/_dlap_loader_char_valid.npz", "/tmp/_dlap_loader_macro_train.npz")   # no macro_idx
    blob.update({"train/macro_all": full.macro_features, "train/mean_all": full.mean_macro, "train/std_all": full.std_macro})
    # a series longer than numpy's pairwise block (128) and than two of them: the recursive split of the column sums
    long_raw = (rng.standard_normal((300, 6)) * rng.uniform(1e-2, 1e2, (1, 6)) + rng.uniform(-20, 20, (1, 6))).astype(np.float32)
    np.savez("/tmp/_dlap_loader_macro_long.npz", data=long_raw)
    lg = AssetPricingDataset("/tmp/_dlap_loader_char_valid.npz", "/tmp/_dlap_loader_macro_long.npz", macro_idx=[5, 1, 2])
    la = AssetPricingDataset("/tmp/_dlap_loader_char_valid.npz", "/tmp/_dlap_loader_macro_long.npz")
    blob.update({"long/macro_raw": long_raw, "long/idx": np.array([5, 1, 2]), "long/macro": lg.macro_features,
                 "long/mean": lg.mean_macro, "long/std": lg.std_macro, "long/macro_all": la.macro_features,
                 "long/mean_all": la.mean_macro, "long/std_all": la.std_macro})
    small = create_small_sample(tr, n_periods=12, n_stocks=40)
    for k, v in small.items():
        blob["small/" + k] = v.numpy()
    # fp64 raw file: the loader rounds to fp32 first (src/data_loader.py:45-46)
    raw64 = files["valid"][0].astype(np.float64) * (1.0 + 1e-9)
    raw64[files["valid"][0] == np.float32(-99.99)] = -99.99
    np.savez("/tmp/_dlap_loader_char64.npz", data=raw64)
    d64 = AssetPricingDataset("/tmp/_dlap_loader_char64.npz")
    blob.update({"f64/raw": raw64, "f64/mask": d64.mask, "f64/returns": d64.returns, "f64/features": d64.individual_features})
    np.savez_compressed(os.path.join(HERE, "loader.npz"), **blob)
    print("loader: train valid", int(tr.mask.sum()), "of", tr.mask.size, "| mean/std dtype", tr.mean_macro.dtype,
          "| min std", float(tr.std_macro.min()))


if __name__ == "__main__":
    torch.set_num_threads(1)        # bit-reproducible regardless of the host
    if sys.argv[1:] == ["variants"]:
        variants()
    elif sys.argv[1:] == ["loader"]:
        loader()
    elif sys.argv[1:] == ["residual"]:
        residual_case()
    else:
        loader()
        residual_case()
        rand_small()
        shipped()
        variants()
