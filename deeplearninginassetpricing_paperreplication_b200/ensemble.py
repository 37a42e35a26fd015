"""BASELINE.json configs[3]: the 9-model ensemble of ``src/evaluate_ensemble.py`` with one model per GPU.

The reference loads the checkpoints one after the other on one device, collects ``get_weights(normalized=True)`` of
every model as numpy arrays, averages them, renormalises each month to unit absolute weight and reports the Sharpe
ratio of the (negated) portfolio (``src/evaluate_ensemble.py:112-171``).  The members are independent, so here every
rank of a process group evaluates the members ``rank, rank + world, ...`` on its own GPU (replicas: nothing is
exchanged while they run), sums their weights on the device, and ONE sum-all-reduce per split assembles the average.
Without a process group the same code runs all members on one GPU.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from .model import AssetPricingGAN


def _sharpe(pr: torch.Tensor) -> float:
    """compute_sharpe of src/evaluate_ensemble.py:46-50 (numpy: population standard deviation)."""
    sd = pr.std(unbiased=False)
    return 0.0 if float(sd) < 1e-8 else float(pr.mean() / sd)


def ensemble_weights(members: Sequence[dict], config: dict, panels: Dict[str, dict], device,
                     group=None) -> Dict[str, dict]:
    """members: state_dicts (``best_model_sharpe.pt`` of each run); config: their common model config (``config.json``);
    panels: {'train': batch, 'valid': ..., 'test': ...} as ``AssetPricingDataset.get_full_batch()`` returns them.
    -> per split: ``weights`` (averaged and renormalised, [T,N] on the device), ``sharpe`` (of the negated portfolio,
    the paper's sign convention, src/evaluate_ensemble.py:169-171) and ``member_sharpe`` (this rank's members)."""
    import torch.distributed as dist
    device = torch.device(device)
    rank = dist.get_rank(group) if group is not None else 0
    world = dist.get_world_size(group) if group is not None else 1
    model = AssetPricingGAN(config).to(device).eval()
    on_dev = {name: {k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in b.items()}
              for name, b in panels.items()}
    sums = {name: torch.zeros_like(b['returns']) for name, b in on_dev.items()}
    member_sharpe: Dict[str, List[float]] = {name: [] for name in on_dev}
    with torch.no_grad():
        for i in range(rank, len(members), world):
            model.load_state_dict(members[i])
            for name, b in on_dev.items():
                w, _ = model.get_weights(b.get('macro_features'), b['individual_features'], b['mask'], normalized=True)
                sums[name] += w
                member_sharpe[name].append(_sharpe((w * b['returns'] * b['mask']).sum(dim=1)))
        out = {}
        for name, b in on_dev.items():
            if group is not None:
                dist.all_reduce(sums[name], group=group)
            w = sums[name] / float(len(members))
            m = b['mask'].to(w.dtype)
            s = (w.abs() * m).sum(dim=1, keepdim=True)
            w = torch.where(s > 1e-8, w / s.clamp_min(1e-30), w)          # normalize_weights, :142-149
            pr = (w * b['returns'] * m).sum(dim=1)
            out[name] = {"weights": w, "sharpe": _sharpe(-pr), "member_sharpe": member_sharpe[name]}
    return out
