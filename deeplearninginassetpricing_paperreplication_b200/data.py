"""Device-side loader: the reference's ``src/data_loader.py`` with the array work done on the GPU.

``AssetPricingDataset`` has the reference's constructor, attributes and methods (src/data_loader.py:11-153); the
difference is where the arrays live.  The ``.npz`` file is read and decompressed by numpy (that is file I/O, not
the path), its raw ``data`` array goes into page-locked host memory, and ONE kernel (``dlap_panel_ingest``) reads it
over the bus and leaves the step's inputs in HBM: the mask of src/data_loader.py:50-57, the zero-filled returns and
characteristics of :60-65.  The macro series is z-scored by ``dlap_macro_zscore`` in numpy's own summation order
(:76-86).  All of it is bit-identical to the reference's arrays (tests/test_ingest_gpu.py).  ``get_full_batch()``
returns CUDA tensors, so the ``batch[k].to(device)`` of every epoch (src/train.py:69-75, :126-130) is a no-op and the
valid-cell index of the tensor-core kernels is built once, here.

No CPU fallback: a CUDA device is required.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from . import functional as Fn

MISSING_VALUE = -99.99                                   # src/data_loader.py:52
VALID_ABOVE = float(np.float32(MISSING_VALUE + 1))       # the comparison of :53-56 runs in fp32


def _ptr(t):
    return t.data_ptr() if t is not None else None


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def ingest_panel(raw, device="cuda", valid_above: float = VALID_ABOVE):
    """raw: the ``data`` array of a characteristics file, [T, N, F+1] fp32 or fp64 -- a numpy array, a CPU tensor
    (pinned or not) or a CUDA tensor.  Returns (returns [T,N] fp32, individual_features [T,N,F] fp32, mask [T,N] bool)
    on `device`, equal bit for bit to what src/data_loader.py:45-65 computes."""
    device = torch.device(device)
    if device.type != "cuda":
        raise ValueError("ingest_panel needs a CUDA device: there is no CPU fallback")
    if isinstance(raw, np.ndarray):
        if raw.dtype not in (np.float32, np.float64):
            raw = raw.astype(np.float32)                 # integer files: the reference's astype, on the host
        raw = torch.from_numpy(np.ascontiguousarray(raw))
    if raw.dim() != 3 or raw.shape[2] < 2:
        raise ValueError("raw must be [T, N, F+1] with the return in column 0")
    if raw.dtype not in (torch.float32, torch.float64):
        raw = raw.float()
    raw = raw.contiguous()
    if raw.device.type == "cpu" and not raw.is_pinned():
        raw = raw.pin_memory()                           # the kernel reads host memory directly: it has to be page-locked
    elif raw.device.type == "cuda" and raw.device != device:
        raw = raw.to(device)
    T, N, W = raw.shape
    F = W - 1
    with torch.cuda.device(device):
        x = torch.empty(T, N, F, dtype=torch.float32, device=device)
        R = torch.empty(T, N, dtype=torch.float32, device=device)
        mask = torch.empty(T, N, dtype=torch.uint8, device=device)
        if T * N:
            ws = Fn.PlanOnly(T, N, F, device)
            _lib.call("dlap_panel_ingest", ws.pp, _ptr(raw), int(raw.dtype == torch.float64), float(valid_above),
                      _ptr(x), _ptr(R), _ptr(mask), 0, _stream(device))
            if raw.device.type == "cpu":
                torch.cuda.current_stream(device).synchronize()      # `raw` (host) must outlive the kernel that reads it
    return R, x, mask.view(torch.bool)


def zscore_macro(macro_raw, mean=None, std=None, device="cuda", pairwise=False):
    """src/data_loader.py:76-86 on the device.  macro_raw [T,M]; mean/std [1,M] (or [M]) from the training file, or
    None to compute them.  pairwise: sum in the order numpy uses after the ``macro_idx`` selection (see
    dlap_macro_zscore).  Returns (normalized [T,M], mean [1,M], std [1,M]) as fp32 CUDA tensors."""
    device = torch.device(device)
    if device.type != "cuda":
        raise ValueError("zscore_macro needs a CUDA device: there is no CPU fallback")
    m = torch.as_tensor(np.ascontiguousarray(macro_raw) if isinstance(macro_raw, np.ndarray) else macro_raw)
    m = m.to(device=device, dtype=torch.float32).contiguous()
    T, M = m.shape
    given = mean is not None
    if given:
        mean_d = torch.as_tensor(mean).to(device=device, dtype=torch.float32).reshape(1, M).contiguous().clone()
        std_d = torch.as_tensor(std).to(device=device, dtype=torch.float32).reshape(1, M).contiguous().clone()
    else:
        mean_d = torch.empty(1, M, dtype=torch.float32, device=device)
        std_d = torch.empty(1, M, dtype=torch.float32, device=device)
    out = torch.empty_like(m)
    with torch.cuda.device(device):
        _lib.call("dlap_macro_zscore", _ptr(m), T, M, _ptr(mean_d), _ptr(std_d), int(given), int(pairwise),
                  _ptr(out), _stream(device))
    return out, mean_d, std_d


class AssetPricingDataset(torch.utils.data.Dataset):
    """Mirror of the reference's ``AssetPricingDataset`` (src/data_loader.py:11-153): same constructor arguments
    (+ ``device``), same attributes, but ``returns``, ``individual_features``, ``mask``, ``macro_features``,
    ``mean_macro``, ``std_macro`` are CUDA tensors produced by the two loader kernels."""

    def __init__(self, path_individual_feature: str, path_macro_feature: str = None, macro_idx: list = None,
                 mean_macro=None, std_macro=None, normalize_macro: bool = True, device="cuda"):
        self.device = torch.device(device)
        individual_data = np.load(path_individual_feature)
        data = individual_data["data"]                                            # [T, N, F+1]
        self.returns, self.individual_features, self.mask = ingest_panel(data, self.device)
        self.dates = individual_data.get("date", np.arange(data.shape[0]))
        self.variable_names = individual_data.get("variable", None)

        if path_macro_feature is not None:
            macro = np.load(path_macro_feature)["data"].astype(np.float32)        # [T, M]: a few hundred kilobytes
            if macro_idx is not None:
                macro = macro[:, macro_idx]                                       # src/data_loader.py:72-73
            if normalize_macro:
                self.macro_features, self.mean_macro, self.std_macro = zscore_macro(
                    macro, mean_macro, std_macro, self.device, pairwise=macro_idx is not None)
            else:
                self.macro_features = torch.from_numpy(np.ascontiguousarray(macro)).to(self.device)
                self.mean_macro = self.std_macro = None
        else:
            self.macro_features = self.mean_macro = self.std_macro = None

        self.T, self.N = self.returns.shape
        self.individual_feature_dim = self.individual_features.shape[2]
        self.macro_feature_dim = self.macro_features.shape[1] if self.macro_features is not None else 0
        if self.T * self.N and self.individual_feature_dim <= Fn.TC_MAX_F:
            with torch.cuda.device(self.device):   # the tensor-core kernels' valid-cell list: once per file, not per epoch
                Fn.panel_index(Fn.PlanOnly(self.T, self.N, self.individual_feature_dim, self.device), self.mask)

    def __len__(self):
        return self.T

    def __getitem__(self, idx):
        item = {"individual_features": self.individual_features[idx], "returns": self.returns[idx],
                "mask": self.mask[idx]}
        if self.macro_features is not None:
            item["macro_features"] = self.macro_features[idx]
        return item

    def get_full_batch(self):
        """src/data_loader.py:125-146; the tensors are already on the device."""
        batch = {"individual_features": self.individual_features, "returns": self.returns, "mask": self.mask}
        if self.macro_features is not None:
            batch["macro_features"] = self.macro_features
        return batch

    def get_macro_stats(self):
        return self.mean_macro, self.std_macro

    def get_date_count_list(self):
        """Valid stocks per month as a float32 numpy array, like the reference (src/data_loader.py:152-154)."""
        return self.mask.sum(dim=1).to(torch.float32).cpu().numpy()


def create_data_loaders(train_individual_path: str, train_macro_path: str, valid_individual_path: str,
                        valid_macro_path: str, test_individual_path: str = None, test_macro_path: str = None,
                        macro_idx: list = None, batch_size: int = None, device="cuda"):
    """src/data_loader.py:157-208: validation and test files are normalised with the training file's statistics."""
    train = AssetPricingDataset(train_individual_path, train_macro_path, macro_idx=macro_idx, device=device)
    mean_macro, std_macro = train.get_macro_stats()
    valid = AssetPricingDataset(valid_individual_path, valid_macro_path, macro_idx=macro_idx, mean_macro=mean_macro,
                                std_macro=std_macro, device=device)
    test = None
    if test_individual_path is not None:
        test = AssetPricingDataset(test_individual_path, test_macro_path, macro_idx=macro_idx, mean_macro=mean_macro,
                                   std_macro=std_macro, device=device)
    return train, valid, test


def create_small_sample(dataset: AssetPricingDataset, n_periods: int = 50, n_stocks: int = 100):
    """src/data_loader.py:212-237: the first n_periods months of the n_stocks stocks with the most valid months.
    The ranking is numpy's argsort on the host (N integers), so ties fall exactly as in the reference."""
    T = min(n_periods, dataset.T)
    N = min(n_stocks, dataset.N)
    valid_counts = dataset.mask.sum(dim=0).cpu().numpy()
    top = torch.from_numpy(np.argsort(valid_counts)[-N:].copy()).to(dataset.device)
    sample = {"individual_features": dataset.individual_features[:T].index_select(1, top).contiguous(),
              "returns": dataset.returns[:T].index_select(1, top).contiguous(),
              "mask": dataset.mask[:T].index_select(1, top).contiguous()}
    if dataset.macro_features is not None:
        sample["macro_features"] = dataset.macro_features[:T]
    return sample
