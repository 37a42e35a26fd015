"""CPU-side checks: the C-ABI library builds, loads and exports every symbol include/dlap.h
declares; the Python mirror of the reference's class API has the same surface (constructor dict,
state_dict keys/shapes, default init) and fails loudly -- never falls back -- off the fast path."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from _dlap_testlib import ROOT, golden_params, load_golden
from deeplearninginassetpricing_paperreplication_b200 import (AssetPricingGAN, MacroLSTM, MomentNetwork,
                                                              SDFNetwork, _lib)

HEADER = os.path.join(ROOT, "include", "dlap.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dlap_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    path = _lib.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    names = _declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dlap.h but not exported"
    assert set(_lib.EXPORTS) == set(names)
    assert _lib.load().dlap_abi_version() == 5
    assert _lib.load().dlap_error_string(-2).decode().startswith("dimension")


def test_ctypes_signatures_have_the_arity_of_the_header():
    """Every entry point is bound with as many arguments as include/dlap.h declares (a mismatch would pass garbage
    pointers to the kernels instead of failing)."""
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    decl = {m.group(1): m.group(2) for m in re.finditer(r"\bint\s+(dlap_[a-z0-9_]+)\s*\(([^;]*?)\)\s*;", src, flags=re.S)}
    for name, sig in _lib.SIGNATURES.items():
        assert name in decl, name
        args = decl[name].strip()
        n = 0 if args in ("", "void") else args.count(",") + 1
        assert n == len(sig), (name, n, len(sig))


def test_plan_without_device_reports_error():
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    dims = _lib.DlapDims(240, 10000, 46, 178, 4)
    plan = _lib.DlapPlan()
    rc = _lib.load().dlap_make_plan(ctypes.byref(dims), 0, ctypes.byref(plan))
    assert rc != 0
    bad = _lib.DlapDims(240, 10000, 99, 178, 4)
    assert _lib.load().dlap_make_plan(ctypes.byref(bad), 0, ctypes.byref(plan)) == -2


@pytest.mark.parametrize("name", ["rand_small", "shipped_sub"])
def test_state_dict_surface_and_default_init(name):
    g = load_golden(name)
    ref = golden_params(g)
    torch.manual_seed(int(g["seed"]))
    model = AssetPricingGAN({"macro_feature_dim": g["macro"].shape[1], "individual_feature_dim": 46})
    sd = model.state_dict()
    assert list(sd.keys()) == list(ref.keys())          # same 12 keys, same order
    for k, v in sd.items():
        assert tuple(v.shape) == ref[k].shape, k
        assert np.array_equal(v.numpy(), ref[k]), k     # same seed -> same initial weights
    model.load_state_dict({k: torch.from_numpy(v) for k, v in ref.items()})


def test_param_counts_match_notebook():
    model = AssetPricingGAN({"macro_feature_dim": 178, "individual_feature_dim": 46})
    n = lambda m: sum(p.numel() for p in m.parameters())
    assert (n(model), n(model.sdf_net), n(model.moment_net)) == (12233, 10433, 1800)


def test_no_cpu_fallback():
    model = AssetPricingGAN({"macro_feature_dim": 8, "individual_feature_dim": 46})
    T, N = 4, 5
    x = torch.zeros(T, N, 46); R = torch.zeros(T, N); m = torch.ones(T, N, dtype=torch.bool)
    macro = torch.zeros(T, 8)
    with pytest.raises(ValueError, match="CUDA"):
        model(macro, x, R, m)
    with pytest.raises(ValueError, match="CUDA"):
        model.get_weights(macro, x, m, normalized=True)


def test_constructors_accept_reference_configs():
    # train.py's full config dict, including the keys the model never reads (src/train.py:530-561)
    cfg = {'macro_feature_dim': 8, 'individual_feature_dim': 46, 'hidden_dim': [64, 64], 'num_layers': 2,
           'use_rnn': True, 'num_units_rnn': [4], 'num_layers_rnn': 1, 'cell_type_rnn': 'lstm',
           'hidden_dim_moment': [], 'num_layers_moment': 0, 'num_condition_moment': 8,
           'use_rnn_moment': True, 'num_units_rnn_moment': [32], 'num_layers_rnn_moment': 1,
           'cell_type_rnn_moment': 'lstm', 'dropout': 0.05, 'normalize_w': True, 'weighted_loss': True,
           'residual_loss_factor': 0.0}
    m = AssetPricingGAN(cfg)
    assert m.config is cfg and isinstance(m.sdf_net, SDFNetwork) and isinstance(m.moment_net, MomentNetwork)
    assert isinstance(m.sdf_net.macro_lstm, MacroLSTM)
    assert AssetPricingGAN({'individual_feature_dim': 46, 'num_units_rnn': 4}).sdf_net.macro_lstm is None
    # the notebook variant (notebooks/demo_synthetic.ipynb cell 10) constructs with the same tree
    v = AssetPricingGAN({'macro_feature_dim': 8, 'individual_feature_dim': 46, 'use_rnn': False,
                         'hidden_dim': [64, 32], 'hidden_dim_moment': [32], 'num_condition_moment': 4})
    assert sum(p.numel() for p in v.parameters()) == 7525


def test_stepoutputs_lazy_moments():
    from deeplearninginassetpricing_paperreplication_b200.model import StepOutputs
    calls = []
    out = StepOutputs({'loss': 1.0}, moments_fn=lambda: calls.append(1) or "Q")
    assert out.get('loss_unconditional', 7) == 7 and out['loss'] == 1.0 and not calls
    assert out['moments'] == "Q" and out['moments'] == "Q" and len(calls) == 1
