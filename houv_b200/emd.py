"""Drop-in for /root/reference/utils/metrics/EMD/emd_module.py (`emdFunction`, `emdModule`) and for
`calc_emd` (registration/model_utils_completion.py, completion/model_utils.py): the auction-based EMD
approximation through houv_emd_forward / houv_emd_backward (csrc/emd.cu), every iteration inside one launch.

Same inputs and outputs: xyz1 (prediction), xyz2 (ground truth) [B,n,3] normalised to [0,1], `eps`, `iters`
-> dist [B,n] float32 (squared distance to the matched point), assignment [B,n] int32; gradient for xyz1 only
(emd_module.py:16, :77-86).  Differences, all host-side: no JIT compilation at import, no scratch tensors
(emd_module.py:57-68; the state lives in shared memory), n need not be a multiple of 1024 nor B <= 512, and bad
inputs raise RuntimeError instead of printf + return code.
"""
import torch
from torch import nn
from torch.autograd import Function

from . import _lib


def emd_forward(xyz1, xyz2, eps, iters, want_iters=False):
    _lib.require_cuda_f32(xyz1, "xyz1", 3)
    _lib.require_cuda_f32(xyz2, "xyz2", 3)
    if xyz1.shape != xyz2.shape:
        raise RuntimeError("emd: the two point clouds must have the same shape, got %s and %s"
                           % (tuple(xyz1.shape), tuple(xyz2.shape)))
    if int(iters) <= 0 or not float(eps) >= 0.0:
        raise RuntimeError("emd: iters must be positive and eps non-negative")
    xyz1, xyz2 = xyz1.contiguous(), xyz2.contiguous()
    B, n, _ = xyz1.shape
    dev = xyz1.device
    dist = torch.empty(B, n, device=dev)  # every element is written by the kernel
    assignment = torch.empty(B, n, device=dev, dtype=torch.int32)
    used = torch.empty(B, device=dev, dtype=torch.int32) if want_iters else None
    lib = _lib.load()
    need = lib.houv_emd_workspace_bytes(B, n)
    ws = torch.empty(need, dtype=torch.uint8, device=dev) if need else None
    with torch.cuda.device(dev):
        rc = lib.houv_emd_forward(xyz1.data_ptr(), xyz2.data_ptr(), B, n, float(eps), int(iters),
                                  dist.data_ptr(), assignment.data_ptr(),
                                  used.data_ptr() if want_iters else 0,
                                  ws.data_ptr() if need else 0, need, _lib.stream_ptr())
    _lib.check(rc, "houv_emd_forward")
    _lib.count_launches(1)
    return (dist, assignment, used) if want_iters else (dist, assignment)


class emdFunction(Function):
    """emd_module.py:40-86."""

    @staticmethod
    def forward(ctx, xyz1, xyz2, eps, iters):
        xyz1 = xyz1.contiguous().float()
        xyz2 = xyz2.contiguous().float()
        dist, assignment = emd_forward(xyz1, xyz2, eps, iters)
        ctx.save_for_backward(xyz1, xyz2, assignment)
        ctx.mark_non_differentiable(assignment)
        return dist, assignment

    @staticmethod
    def backward(ctx, graddist, gradidx):
        xyz1, xyz2, assignment = ctx.saved_tensors
        graddist = graddist.contiguous()
        gradxyz1 = torch.zeros_like(xyz1)
        gradxyz2 = torch.zeros_like(xyz2)
        B, n, _ = xyz1.shape
        with torch.cuda.device(xyz1.device):
            rc = _lib.load().houv_emd_backward(xyz1.data_ptr(), xyz2.data_ptr(), graddist.data_ptr(),
                                               assignment.data_ptr(), B, n, gradxyz1.data_ptr(),
                                               _lib.stream_ptr())
        _lib.check(rc, "houv_emd_backward")
        _lib.count_launches(1)
        return gradxyz1, gradxyz2, None, None


class emdModule(nn.Module):
    """emd_module.py:88-93."""

    def __init__(self):
        super(emdModule, self).__init__()

    def forward(self, input1, input2, eps, iters):
        return emdFunction.apply(input1, input2, eps, iters)


emd = emdModule  # utils/metrics/EMD/__init__.py


def calc_emd(output, gt, eps=0.005, iterations=50):
    """registration/model_utils_completion.py `calc_emd`: sqrt of the matched squared distances, mean per sample."""
    dist, _ = emdModule()(output, gt, eps, iterations)
    return torch.sqrt(dist).mean(1)
