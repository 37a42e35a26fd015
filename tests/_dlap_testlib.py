"""Helpers shared by the test modules (fixture loading, tolerances)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
PHASES = ("unconditional", "moment", "conditional")


def load_golden(name):
    """One fixture written by tests/golden/make_golden.py (outputs of the reference itself)."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: z[k] for k in z.files}


def golden_params(g):
    return {k[len("param/"):]: v for k, v in g.items() if k.startswith("param/")}


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
