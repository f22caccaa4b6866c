"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by houv_b200/).

ctypes front of oracle/emd_oracle.c, the CPU restatement of the reference's EMD auction
(/root/reference/utils/metrics/EMD/emd_cuda.cu:95-266), and a loader for the reference's own CUDA
extension built unmodified for sm_100a (`make -C oracle ref-emd` -> oracle/_ref/emd.so), with the buffer
allocation of emd_module.py:45-70 restated on device tensors.
"""
import ctypes
import importlib.util
import os

import numpy as np

from . import chamfer as _oc

_f = ctypes.POINTER(ctypes.c_float)
_i = ctypes.POINTER(ctypes.c_int)
_REF_EMD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "emd.so")
_mod = None


def emd_forward(xyz1, xyz2, eps, iters):
    """xyz1, xyz2 [B,n,3] float32 -> dist [B,n] float32, assignment [B,n] int32, iterations used [B]."""
    x1 = np.ascontiguousarray(xyz1, np.float32)
    x2 = np.ascontiguousarray(xyz2, np.float32)
    B, n, _ = x1.shape
    assert x2.shape == x1.shape
    dist = np.zeros((B, n), np.float32)
    asg = np.zeros((B, n), np.int32)
    used = np.zeros(B, np.int32)
    rc = _oc.oracle_lib().emd_oracle_forward(
        x1.ctypes.data_as(_f), x2.ctypes.data_as(_f), ctypes.c_int(B), ctypes.c_int(n),
        ctypes.c_float(eps), ctypes.c_int(iters), dist.ctypes.data_as(_f), asg.ctypes.data_as(_i),
        None, used.ctypes.data_as(_i))
    assert rc == 0
    return dist, asg, used


def emd_backward(xyz1, xyz2, graddist, assignment):
    x1 = np.ascontiguousarray(xyz1, np.float32)
    x2 = np.ascontiguousarray(xyz2, np.float32)
    gd = np.ascontiguousarray(graddist, np.float32)
    a = np.ascontiguousarray(assignment, np.int32)
    B, n, _ = x1.shape
    g = np.zeros_like(x1)
    _oc.oracle_lib().emd_oracle_backward(x1.ctypes.data_as(_f), x2.ctypes.data_as(_f),
                                         gd.ctypes.data_as(_f), a.ctypes.data_as(_i),
                                         ctypes.c_int(B), ctypes.c_int(n), g.ctypes.data_as(_f))
    return g


def ref_available():
    return os.path.exists(_REF_EMD)


def ref_load():
    global _mod
    if _mod is None:
        import torch  # noqa: F401
        spec = importlib.util.spec_from_file_location("emd", _REF_EMD)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


def ref_emd_forward(xyz1, xyz2, eps, iters):
    """emd_module.py:45-70 on device tensors through the reference extension -> (dist, assignment)."""
    import torch
    B, n, _ = xyz1.shape
    dev = xyz1.device
    z = lambda *s, dt=torch.float32: torch.zeros(*s, device=dev, dtype=dt)
    dist = z(B, n)
    assignment = z(B, n, dt=torch.int32) - 1
    assignment_inv = z(B, n, dt=torch.int32) - 1
    price = z(B, n)
    bid = z(B, n, dt=torch.int32)
    bid_increments = z(B, n)
    max_increments = z(B, n)
    unass_idx = z(B * n, dt=torch.int32)
    max_idx = z(B * n, dt=torch.int32)
    unass_cnt, unass_cnt_sum, cnt_tmp = (z(512, dt=torch.int32) for _ in range(3))
    ref_load().forward(xyz1.contiguous(), xyz2.contiguous(), dist, assignment, price, assignment_inv, bid,
                       bid_increments, max_increments, unass_idx, unass_cnt, unass_cnt_sum, cnt_tmp,
                       max_idx, eps, iters)
    return dist, assignment
