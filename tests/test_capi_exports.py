"""CPU: the C-ABI library builds, loads, and exports every symbol that
include/houv_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

from houv_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "houv_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(houv_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_header_symbols():
    path = build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    names = _declared()
    assert "houv_chamfer3d_forward" in names and "houv_chamfer3d_backward" in names
    for n in names:
        assert hasattr(lib, n), "missing export " + n


def test_binding_table_matches_header():
    assert sorted(_lib.EXPORTS) == _declared()


def test_version_string():
    assert _lib.load().houv_version().decode().startswith("houv_b200")


def test_no_cpu_fallback():
    import pytest
    import torch
    from houv_b200.chamfer import chamfer_3DDist
    with pytest.raises(RuntimeError):
        chamfer_3DDist()(torch.rand(1, 8, 3), torch.rand(1, 8, 3))
