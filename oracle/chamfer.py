"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by houv_b200/).

ctypes front-ends, on numpy arrays, for
  * oracle/_build/libhouv_oracle.so  -- our C restatement (chamfer_oracle.c)
  * oracle/_ref/libref_chamfer.so    -- the reference's own kernel bodies built
                                        for the host (oracle/Makefile `ref`)
Both follow chamfer_cuda_forward / chamfer_cuda_backward
(/root/reference/utils/metrics/CD/chamfer3D/chamfer3D.cu:136-154, 176-195).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(_HERE, "_build", "libhouv_oracle.so")
_REF_SO = os.path.join(_HERE, "_ref", "libref_chamfer.so")

_f = ctypes.POINTER(ctypes.c_float)
_d = ctypes.POINTER(ctypes.c_double)
_i = ctypes.POINTER(ctypes.c_int)


def build(ref=True):
    """Compile the checkers (gcc only). `ref` also rebuilds oracle/_ref when the
    reference tree is present; otherwise the prebuilt file is kept."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"] + (["ref"] if ref else []))


def _load(path):
    if not os.path.exists(path):
        build(ref=(path == _REF_SO))
    return ctypes.CDLL(path)


_libs = {}


def oracle_lib():
    if "o" not in _libs:
        _libs["o"] = _load(_ORACLE_SO)
    return _libs["o"]


def ref_available():
    return os.path.exists(_REF_SO) or os.path.exists(
        "/root/reference/utils/metrics/CD/chamfer3D/chamfer3D.cu")


def ref_lib():
    if "r" not in _libs:
        _libs["r"] = _load(_REF_SO)
    return _libs["r"]


def set_threads(t):
    oracle_lib().oracle_set_threads(ctypes.c_int(int(t)))


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def _p(a, t):
    return a.ctypes.data_as(t)


def nn_forward(query, target):
    """Directed search (NmDistanceKernel, chamfer3D.cu:12-134): query[B,N,3],
    target[B,M,3] -> dist[B,N] f32, idx[B,N] i32 (first minimum)."""
    q, t = _c(query, np.float32), _c(target, np.float32)
    b, n, _ = q.shape
    m = t.shape[1]
    dist = np.zeros((b, n), np.float32)
    idx = np.zeros((b, n), np.int32)
    oracle_lib().oracle_nn_forward(b, n, _p(q, _f), m, _p(t, _f), _p(dist, _f), _p(idx, _i))
    return dist, idx


def _forward(fn, xyz1, xyz2):
    x1, x2 = _c(xyz1, np.float32), _c(xyz2, np.float32)
    b, n, _ = x1.shape
    m = x2.shape[1]
    d1, d2 = np.zeros((b, n), np.float32), np.zeros((b, m), np.float32)
    i1, i2 = np.zeros((b, n), np.int32), np.zeros((b, m), np.int32)
    fn(b, n, _p(x1, _f), m, _p(x2, _f), _p(d1, _f), _p(i1, _i), _p(d2, _f), _p(i2, _i))
    return d1, d2, i1, i2


def chamfer_forward(xyz1, xyz2):
    """-> dist1[B,N], dist2[B,M], idx1[B,N], idx2[B,M] (restatement)."""
    return _forward(oracle_lib().oracle_chamfer_forward, xyz1, xyz2)


def ref_chamfer_forward(xyz1, xyz2):
    """Same, through the reference's own kernel source built for the host."""
    return _forward(ref_lib().ref_chamfer_forward, xyz1, xyz2)


def _backward(fn, xyz1, xyz2, gd1, gd2, idx1, idx2, dt, pt):
    x1, x2 = _c(xyz1, np.float32), _c(xyz2, np.float32)
    g1, g2 = _c(gd1, np.float32), _c(gd2, np.float32)
    i1, i2 = _c(idx1, np.int32), _c(idx2, np.int32)
    b, n, _ = x1.shape
    m = x2.shape[1]
    o1, o2 = np.zeros((b, n, 3), dt), np.zeros((b, m, 3), dt)
    fn(b, n, _p(x1, _f), m, _p(x2, _f), _p(g1, _f), _p(i1, _i), _p(g2, _f), _p(i2, _i),
       _p(o1, pt), _p(o2, pt))
    return o1, o2


def chamfer_backward(xyz1, xyz2, gd1, gd2, idx1, idx2, f64=False):
    """-> gradxyz1[B,N,3], gradxyz2[B,M,3]; float32 in index order, or float64."""
    L = oracle_lib()
    if f64:
        return _backward(L.oracle_chamfer_backward_f64, xyz1, xyz2, gd1, gd2, idx1, idx2,
                         np.float64, _d)
    return _backward(L.oracle_chamfer_backward_f32, xyz1, xyz2, gd1, gd2, idx1, idx2,
                     np.float32, _f)


def ref_chamfer_backward(xyz1, xyz2, gd1, gd2, idx1, idx2):
    return _backward(ref_lib().ref_chamfer_backward, xyz1, xyz2, gd1, gd2, idx1, idx2,
                     np.float32, _f)
