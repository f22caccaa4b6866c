"""ctypes binding of the C ABI (include/houv_b200.h).  No CPU fallback: if the
library is missing it is built with nvcc, and if that fails the import of any
operator raises."""
import ctypes
import os

from . import build as _build

_LIB = None

_vp = ctypes.c_void_p
_ci = ctypes.c_int
_cf = ctypes.c_float

_fp = ctypes.c_void_p  # device pointers travel as integers


class HouvPoseProblem(ctypes.Structure):
    """include/houv_b200.h: struct HouvPoseProblem (same field order)."""
    _fields_ = [("src", _fp), ("tgt", _fp),
                ("P", _ci), ("N", _ci), ("M", _ci), ("K", _ci),
                ("params", _fp), ("adam_m", _fp), ("adam_v", _fp),
                ("params64", _fp), ("adam_m64", _fp), ("adam_v64", _fp),
                ("step0", _ci), ("iters", _ci), ("angle_base", _ci),
                ("lr", ctypes.c_double), ("beta1", ctypes.c_double), ("beta2", ctypes.c_double),
                ("eps", ctypes.c_double),
                ("grad_scale", _cf), ("percent", _cf), ("w3d", _cf), ("wview", _cf),
                ("tran_scale", _cf), ("tran_offset", _cf),
                ("min1_out", _fp), ("loss_out", _fp), ("R_out", _fp), ("T_out", _fp),
                ("tran_s_out", _fp), ("trace", _fp),
                ("row_pair", _fp), ("row_base", _fp),
                ("rows", _ci),
                ("row_scale", _fp)]


# name -> argtypes; every entry point returns int (0 = ok) unless noted
_SIGNATURES = {
    "houv_chamfer3d_forward": [_vp, _vp, _ci, _ci, _ci, _vp, _vp, _vp, _vp, _vp],
    "houv_chamfer3d_forward_static": [_vp, _vp, _ci, _ci, _ci, _vp, _vp, _vp, _vp, _vp],
    "houv_chamfer3d_forward_ws": [_vp, _vp, _ci, _ci, _ci, _vp, _vp, _vp, _vp, _vp,
                                  ctypes.c_size_t, _vp],
    "houv_cd_reduce": [_vp, _vp, _ci, _ci, _ci, _cf, _vp, _vp, _vp, _vp],
    "houv_nn_search": [_vp, _vp, _ci, _ci, _ci, _vp, _vp, _vp],
    "houv_chamfer3d_backward": [_vp, _vp, _vp, _vp, _vp, _vp, _ci, _ci, _ci, _vp, _vp, _vp],
    "houv_emd_forward": [_vp, _vp, _ci, _ci, ctypes.c_float, _ci, _vp, _vp, _vp, _vp, ctypes.c_size_t, _vp],
    "houv_emd_backward": [_vp, _vp, _vp, _vp, _ci, _ci, _vp, _vp],
    "houv_idam_step": [_vp, _vp, _vp, _ci, _vp, _ci, _ci, _ci, _vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "houv_chamfer3d_backward_assign": [_vp, _vp, _vp, _vp, _vp, _vp, _ci, _ci, _ci, _vp, _vp, _vp],
    "houv_chamfer3d_backward_deterministic": [_vp, _vp, _vp, _vp, _vp, _vp, _ci, _ci, _ci, _vp, _vp,
                                              _ci, _vp],
    "houv_kabsch": [_vp, _vp, _vp, _ci, _ci, _vp, _vp, _vp],
    "houv_icp": [_vp, _vp, _ci, _ci, _ci, _vp, _ci, _cf, _ci, ctypes.c_double, ctypes.c_double,
                 _vp, _vp, _vp],
    "houv_set_nn_variant": [_ci],
    "houv_set_pose_variant": [_ci],
    "houv_pose_optimise": [_vp, _vp, _ci, _ci, _ci, _ci, _vp, _vp, _vp, _ci, _ci, _ci,
                           _cf, _cf, _cf, _cf, _cf, _cf, _cf, _cf, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                           _ci, _vp],
    "houv_pose_optimise_ex": [ctypes.POINTER(HouvPoseProblem), _vp],
    "houv_fp32_probe": [_vp, _ci, _ci, ctypes.POINTER(ctypes.c_double), _vp],
}

EXPORTS = tuple(_SIGNATURES) + ("houv_version", "houv_chamfer3d_workspace_bytes",
                                "houv_chamfer3d_static_scratch_bytes", "houv_emd_workspace_bytes")


def lib_path():
    return _build.LIB


def load():
    """Load (building first if needed) libhouv_b200.so."""
    global _LIB
    if _LIB is None:
        path = _build.build()
        lib = ctypes.CDLL(path)
        for name, args in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = _ci
        lib.houv_chamfer3d_static_scratch_bytes.restype = ctypes.c_size_t
        lib.houv_chamfer3d_static_scratch_bytes.argtypes = []
        lib.houv_chamfer3d_workspace_bytes.restype = ctypes.c_size_t
        lib.houv_chamfer3d_workspace_bytes.argtypes = [_ci, _ci, _ci]
        lib.houv_emd_workspace_bytes.restype = ctypes.c_size_t
        lib.houv_emd_workspace_bytes.argtypes = [_ci, _ci]
        lib.houv_version.restype = ctypes.c_char_p
        lib.houv_version.argtypes = []
        _LIB = lib
    return _LIB


class HouvError(RuntimeError):
    pass


LAUNCHES = [0]  # kernel launches issued through the C ABI (bench.py reports them)


def count_launches(n=1):
    LAUNCHES[0] += n


def check(code, what):
    if code != 0:
        raise HouvError("%s failed with code %d" % (what, code))


def stream_ptr():
    """cudaStream_t of torch's current stream as an integer handle."""
    import torch
    return torch.cuda.current_stream().cuda_stream


def require_cuda_f32(t, name, last=None):
    """The reference surfaces dtype/device mismatches as c10::Error from
    .data<float>() (chamfer_cuda.cpp:17-27); we raise RuntimeError up front."""
    import torch
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (houv_b200 has no CPU path)" % name)
    if t.dtype != torch.float32:
        raise RuntimeError("%s must be float32, got %s" % (name, t.dtype))
    if last is not None and (t.dim() != 3 or t.shape[-1] != last):
        raise RuntimeError("%s must be [B, points, %d], got %s" % (name, last, tuple(t.shape)))
