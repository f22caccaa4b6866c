"""Shared helpers for the tests (fixtures live in tests/golden/)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_npz(name):
    return np.load(os.path.join(GOLDEN, name))


def grouped(npz):
    """'case/key' flat npz -> {case: {key: array}}."""
    out = {}
    for k in npz.files:
        c, f = k.split("/")
        out.setdefault(c, {})[f] = npz[k]
    return out
