"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by houv_b200/).

The reference's own CUDA extension, built UNMODIFIED for sm_100a from
/root/reference/utils/metrics/CD/chamfer3D/{chamfer3D.cu,chamfer_cuda.cpp} by
`make -C oracle ref-gpu` into oracle/_ref/chamfer_3D.so (git-ignored, travels to the GPU box).
On a B200 it is the GPU oracle (tests/test_reference_ext_gpu.py) and the GPU baseline
(tools/microbench.py refext).  `chamfer_forward/backward` restate the allocation done by the
reference wrapper (dist_chamfer_3D.py:28-47, 50-64), with the buffers created on the device.
"""
import importlib.util
import os

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "chamfer_3D.so")
_mod = None


def available():
    return os.path.exists(_SO)


def load():
    """The pybind module `chamfer_3D` (chamfer_cuda.cpp:30-33): forward(xyz1, xyz2, dist1, dist2,
    idx1, idx2) and backward(xyz1, xyz2, gradxyz1, gradxyz2, graddist1, graddist2, idx1, idx2)."""
    global _mod
    if _mod is None:
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        spec = importlib.util.spec_from_file_location("chamfer_3D", _SO)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


def chamfer_forward(xyz1, xyz2):
    """dist_chamfer_3D.py:28-47 on device tensors -> (dist1, dist2, idx1, idx2)."""
    import torch
    B, n, _ = xyz1.shape
    m = xyz2.shape[1]
    dev = xyz1.device
    dist1 = torch.zeros(B, n, device=dev)
    dist2 = torch.zeros(B, m, device=dev)
    idx1 = torch.zeros(B, n, device=dev, dtype=torch.int32)
    idx2 = torch.zeros(B, m, device=dev, dtype=torch.int32)
    load().forward(xyz1, xyz2, dist1, dist2, idx1, idx2)
    return dist1, dist2, idx1, idx2


def chamfer_backward(xyz1, xyz2, graddist1, graddist2, idx1, idx2):
    """dist_chamfer_3D.py:50-64 on device tensors -> (gradxyz1, gradxyz2)."""
    import torch
    g1 = torch.zeros_like(xyz1)
    g2 = torch.zeros_like(xyz2)
    load().backward(xyz1, xyz2, g1, g2, graddist1, graddist2, idx1, idx2)
    return g1, g2
