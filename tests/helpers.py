"""Shared helpers for the tests: golden-fixture access."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_npz(name):
    z = np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    meta = json.loads(str(z["__meta__"]))
    return z, meta


def relerr(x, ref):
    x = np.asarray(x)
    ref = np.asarray(ref)
    scale = np.maximum(np.abs(ref), 1e-300)
    return float(np.max(np.abs(x - ref) / scale)) if ref.size else 0.0


def phase_err(x, ref):
    d = np.asarray(x) - np.asarray(ref)
    d = (d + np.pi) % (2 * np.pi) - np.pi
    return float(np.max(np.abs(d))) if d.size else 0.0
