"""configs[3]: ensemble of models, weights averaged on the device, against the reference's numpy procedure
(src/evaluate_ensemble.py:112-171) on the shipped validation / test splits."""
import numpy as np
import pytest
import torch

from deeplearninginassetpricing_paperreplication_b200 import AssetPricingGAN
from deeplearninginassetpricing_paperreplication_b200.ensemble import ensemble_weights
from test_parity_gpu import dev, make_panel

pytestmark = pytest.mark.gpu


def test_ensemble_average_matches_the_reference_procedure():
    d = dev()
    cfg = {"macro_feature_dim": 8, "individual_feature_dim": 46}
    panels = {}
    for name, T, seed in (("valid", 9, 1), ("test", 14, 2)):
        macro, x, R, mask = make_panel(T, 300, 8, seed=seed)
        panels[name] = {"macro_features": macro, "individual_features": x, "returns": R, "mask": mask}
    members = []
    for i in range(3):
        torch.manual_seed(50 + i)
        members.append(AssetPricingGAN(cfg).state_dict())
    res = ensemble_weights(members, cfg, panels, d)
    model = AssetPricingGAN(cfg).to(d).eval()
    for name, b in panels.items():
        ws = []
        for sd in members:                                   # the reference: one model after the other, numpy
            model.load_state_dict(sd)
            with torch.no_grad():
                w, _ = model.get_weights(b["macro_features"].to(d), b["individual_features"].to(d), b["mask"].to(d),
                                         normalized=True)
            ws.append(w.cpu().numpy())
        avg = np.mean(ws, axis=0)
        m = b["mask"].numpy().astype(float)
        for t in range(avg.shape[0]):
            s = np.abs(avg[t] * m[t]).sum()
            if s > 1e-8:
                avg[t] = avg[t] / s
        pr = (avg * b["returns"].numpy() * m).sum(axis=1)
        sharpe = (-pr).mean() / (-pr).std()
        assert np.abs(res[name]["weights"].cpu().numpy() - avg).max() <= 1e-5 * np.abs(avg).max()
        assert abs(res[name]["sharpe"] - sharpe) <= 1e-4 * abs(sharpe) + 1e-7
        assert len(res[name]["member_sharpe"]) == 3
