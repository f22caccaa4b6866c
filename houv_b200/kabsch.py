"""Drop-in for the reference's Kabsch head
(/root/reference/registration/model_utils.py:213-255 `SVDHead`, and its
duplicates models/idam.py:152-189, models/dcp.py:331-381): same constructor and
forward signature, R[B,3,3] and t[B,3] out.  The per-sample Python loop of
torch.svd + torch.det + host-synchronising `if` (:232-240) is ONE launch of the
warp-per-pair K4 kernel (houv_kabsch)."""
import torch
import torch.nn as nn

from . import _lib


def kabsch(src, src_corr, weights=None):
    """src, src_corr [B,3,N] float32 CUDA; weights [B,1,N] or [B,N] or None -> R[B,3,3], t[B,3]."""
    _lib.require_cuda_f32(src, "src")
    _lib.require_cuda_f32(src_corr, "src_corr")
    if src.dim() != 3 or src.shape[1] != 3 or src.shape != src_corr.shape:
        raise RuntimeError("src and src_corr must both be [B, 3, N], got %s and %s"
                           % (tuple(src.shape), tuple(src_corr.shape)))
    src, src_corr = src.contiguous(), src_corr.contiguous()
    B, _, N = src.shape
    w = None
    if weights is not None:
        _lib.require_cuda_f32(weights, "weights")
        w = weights.reshape(B, N).contiguous()
    R = torch.empty(B, 3, 3, device=src.device)
    t = torch.empty(B, 3, device=src.device)
    with torch.cuda.device(src.device):
        rc = _lib.load().houv_kabsch(src.data_ptr(), src_corr.data_ptr(),
                                     0 if w is None else w.data_ptr(), B, N, R.data_ptr(),
                                     t.data_ptr(), _lib.stream_ptr())
    _lib.check(rc, "houv_kabsch")
    _lib.count_launches(1)
    return R, t


class SVDHead(nn.Module):
    def __init__(self, args=None):
        super(SVDHead, self).__init__()
        if args is not None:  # model_utils.py:216 (kept for state-dict compatibility)
            self.emb_dims = 33 if getattr(args, "use_fpfh", False) else getattr(
                args, "descriptor_size", None)
        self.reflect = nn.Parameter(torch.eye(3), requires_grad=False)
        self.reflect[2, 2] = -1

    def forward(self, src, src_corr, weights=None):
        R, t = kabsch(src, src_corr, weights)
        return R, t.view(src.size(0), 3)


def nearest_neighbor(src, dst):
    """model_utils.py:33-37 semantics (negated squared distance of the nearest dst point and
    its index, for [dims, points] inputs) through the exact K1 search instead of the
    expanded-form matmul + topk."""
    from .chamfer import nn_search
    d, i = nn_search(src.t().unsqueeze(0).contiguous(), dst.t().unsqueeze(0).contiguous())
    return -d[0].unsqueeze(-1), i[0].long().unsqueeze(-1)
