"""Drop-in for the reference's Chamfer operator module
(/root/reference/utils/metrics/CD/chamfer3D/dist_chamfer_3D.py:26-74; exported
as `cd` by utils/metrics/CD/__init__.py:1).

Same class names, call signatures, return order and dtypes
(dist1 f32, dist2 f32, idx1 i32, idx2 i32) and the same tensors saved for
backward.  Differences, all host-side: outputs are allocated on the device
(the reference fills CPU zeros and copies them over, :33-42, :56-60), kernels
run on torch's CURRENT stream under a device guard (the reference uses the
legacy default stream and torch.cuda.set_device, :43), and bad inputs raise
RuntimeError instead of the reference's printf-and-continue.
"""
import torch
from torch import nn
from torch.autograd import Function

from . import _lib


DETERMINISTIC_MAX_POINTS = 2048  # csrc/chamfer_bwd.cu kGMax


class _Chamfer3DModule:
    """Object with the reference pybind module's two entry points
    (chamfer_cuda.cpp:30-33): forward/backward on caller-allocated tensors,
    returning 1 on success as chamfer3D.cu:151,194 do."""

    @staticmethod
    def forward(xyz1, xyz2, dist1, dist2, idx1, idx2):
        B, N, _ = xyz1.shape
        M = xyz2.shape[1]
        lib = _lib.load()
        # scratch for the symmetric kernel when the larger cloud spans several 2048-row CTAs
        # (caller-owned like every other buffer: the C ABI never allocates)
        need = lib.houv_chamfer3d_workspace_bytes(B, N, M)
        ws = torch.empty(need, dtype=torch.uint8, device=xyz1.device) if need else None
        with torch.cuda.device(xyz1.device):
            rc = lib.houv_chamfer3d_forward_ws(
                xyz1.data_ptr(), xyz2.data_ptr(), B, N, M, dist1.data_ptr(), dist2.data_ptr(),
                idx1.data_ptr(), idx2.data_ptr(), ws.data_ptr() if need else 0, need,
                _lib.stream_ptr())
        _lib.check(rc, "houv_chamfer3d_forward_ws")
        _lib.count_launches(1)
        return 1

    @staticmethod
    def backward(xyz1, xyz2, gradxyz1, gradxyz2, graddist1, graddist2, idx1, idx2, assign=False):
        """Accumulates into gradxyz1/gradxyz2 like the reference; assign=True (not in the reference)
        writes them instead, so they may be torch.empty."""
        B, N, _ = xyz1.shape
        M = xyz2.shape[1]
        lib = _lib.load()
        if torch.are_deterministic_algorithms_enabled():
            # torch.use_deterministic_algorithms(True): the gather-side kernel (no float atomics,
            # bit-identical to a sequential run of the reference kernels); clouds up to 2048 points
            if max(N, M) > DETERMINISTIC_MAX_POINTS:
                msg = ("chamfer_3D.backward has no deterministic kernel for clouds above %d points "
                       "(float atomics, like the reference)" % DETERMINISTIC_MAX_POINTS)
                if not torch.is_deterministic_algorithms_warn_only_enabled():
                    raise RuntimeError(msg)
                import warnings
                warnings.warn(msg)
            else:
                with torch.cuda.device(xyz1.device):
                    rc = lib.houv_chamfer3d_backward_deterministic(
                        xyz1.data_ptr(), xyz2.data_ptr(), graddist1.data_ptr(), graddist2.data_ptr(),
                        idx1.data_ptr(), idx2.data_ptr(), B, N, M, gradxyz1.data_ptr(),
                        gradxyz2.data_ptr(), 1 if assign else 0, _lib.stream_ptr())
                _lib.check(rc, "houv_chamfer3d_backward_deterministic")
                _lib.count_launches(1)
                return 1
        fn = lib.houv_chamfer3d_backward_assign if assign else lib.houv_chamfer3d_backward
        with torch.cuda.device(xyz1.device):
            rc = fn(xyz1.data_ptr(), xyz2.data_ptr(), graddist1.data_ptr(), graddist2.data_ptr(),
                    idx1.data_ptr(), idx2.data_ptr(), B, N, M, gradxyz1.data_ptr(),
                    gradxyz2.data_ptr(), _lib.stream_ptr())
        _lib.check(rc, "houv_chamfer3d_backward")
        _lib.count_launches(1)
        return 1


chamfer_3D = _Chamfer3DModule()


def _check_pair(xyz1, xyz2):
    _lib.require_cuda_f32(xyz1, "xyz1", 3)
    _lib.require_cuda_f32(xyz2, "xyz2", 3)
    if xyz1.shape[0] != xyz2.shape[0]:
        raise RuntimeError("batch sizes differ: %d vs %d" % (xyz1.shape[0], xyz2.shape[0]))
    if xyz1.device != xyz2.device:
        raise RuntimeError("xyz1 and xyz2 are on different devices")


class chamfer_3DFunction(Function):
    """dist_chamfer_3D.py:26-64."""

    @staticmethod
    def forward(ctx, xyz1, xyz2):
        _check_pair(xyz1, xyz2)
        xyz1 = xyz1.contiguous()
        xyz2 = xyz2.contiguous()
        batchsize, n, _ = xyz1.size()
        _, m, _ = xyz2.size()
        device = xyz1.device
        dist1 = torch.empty(batchsize, n, device=device, dtype=torch.float32)
        dist2 = torch.empty(batchsize, m, device=device, dtype=torch.float32)
        idx1 = torch.empty(batchsize, n, device=device, dtype=torch.int32)
        idx2 = torch.empty(batchsize, m, device=device, dtype=torch.int32)
        chamfer_3D.forward(xyz1, xyz2, dist1, dist2, idx1, idx2)
        ctx.save_for_backward(xyz1, xyz2, idx1, idx2)
        ctx.mark_non_differentiable(idx1, idx2)
        return dist1, dist2, idx1, idx2

    @staticmethod
    def backward(ctx, graddist1, graddist2, gradidx1, gradidx2):
        xyz1, xyz2, idx1, idx2 = ctx.saved_tensors
        graddist1 = graddist1.contiguous()
        graddist2 = graddist2.contiguous()
        # the reference zero-fills and accumulates (:56-63); the assign entry point writes the
        # gradient in the same launch
        gradxyz1 = torch.empty_like(xyz1)
        gradxyz2 = torch.empty_like(xyz2)
        chamfer_3D.backward(xyz1, xyz2, gradxyz1, gradxyz2, graddist1, graddist2, idx1, idx2,
                            assign=True)
        return gradxyz1, gradxyz2


class chamfer_3DDist(nn.Module):
    """dist_chamfer_3D.py:67-74."""

    def __init__(self):
        super(chamfer_3DDist, self).__init__()

    def forward(self, input1, input2):
        input1 = input1.contiguous()
        input2 = input2.contiguous()
        return chamfer_3DFunction.apply(input1, input2)


cd = chamfer_3DDist  # utils/metrics/CD/__init__.py:1


def nn_search(query, target):
    """One directed search (a single NmDistanceKernel launch, chamfer3D.cu:142):
    query[B,N,3], target[B,M,3] -> dist[B,N] f32, idx[B,N] i32. No autograd."""
    _check_pair(query, target)
    query, target = query.contiguous(), target.contiguous()
    B, N, _ = query.shape
    M = target.shape[1]
    dist = torch.empty(B, N, device=query.device, dtype=torch.float32)
    idx = torch.empty(B, N, device=query.device, dtype=torch.int32)
    with torch.cuda.device(query.device):
        rc = _lib.load().houv_nn_search(query.data_ptr(), target.data_ptr(), B, N, M,
                                        dist.data_ptr(), idx.data_ptr(), _lib.stream_ptr())
    _lib.check(rc, "houv_nn_search")
    _lib.count_launches(1)
    return dist, idx


def fscore(dist1, dist2, threshold=0.0001):
    """utils/metrics/CD/fscore.py:3-16 (plain torch, any device)."""
    precision_1 = torch.mean((dist1 < threshold).float(), dim=1)
    precision_2 = torch.mean((dist2 < threshold).float(), dim=1)
    f = 2 * precision_1 * precision_2 / (precision_1 + precision_2)
    f[torch.isnan(f)] = 0
    return f, precision_1, precision_2
