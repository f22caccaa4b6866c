"""Drop-in for the reference's Kabsch head
(/root/reference/registration/model_utils.py:213-255 `SVDHead` and its duplicate
models/idam.py:152-189): same constructor and forward signature, R[B,3,3] and t[B,3] out.  The
per-sample Python loop of torch.svd + torch.det + host-synchronising `if` (:232-240) is ONE
launch of the warp-per-pair K4 kernel (houv_kabsch: one-sided Jacobi SVD of the 3x3 in float64).

FORWARD ONLY.  Both callers on the path detach the result (idam.py:337-338; train_ICP.py never
differentiates).  DCP's head (models/dcp.py:331-381) has another signature AND is trained by
back-propagation through torch.svd: it is not claimed, and inputs that require grad raise here
instead of silently returning R, t without a graph."""
import torch
import torch.nn as nn

from . import _lib


def kabsch(src, src_corr, weights=None):
    """src, src_corr [B,3,N] float32 CUDA; weights [B,1,N] or [B,N] or None -> R[B,3,3], t[B,3]."""
    _lib.require_cuda_f32(src, "src")
    _lib.require_cuda_f32(src_corr, "src_corr")
    if torch.is_grad_enabled() and any(x is not None and x.requires_grad
                                       for x in (src, src_corr, weights)):
        raise RuntimeError("houv_b200 kabsch / SVDHead is forward-only (no autograd through the "
                           "SVD): detach the inputs as idam.py:337-338 does, or call it under "
                           "torch.no_grad()")
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


def hybrid_point_elimination(weight_logits):
    """idam.py:329-331 in torch ops (autograd-friendly): sigmoid, keep scores >= the row median,
    normalise to sum 1."""
    w = torch.sigmoid(weight_logits)
    w = w * (w >= w.median(-1, keepdim=True)[0]).float()
    return w / (w.sum(-1, keepdim=True) + 1e-8)


def idam_step(src, tgt, corr_idx, weight_logits, R=None, t=None):
    """The tail of one IDAM iteration (idam.py:317-342) in ONE launch (houv_idam_step): gather the
    correspondences, hybrid point elimination, weighted Kabsch, move `src`, compose the running pose.

    src [B,3,N], tgt [B,3,M] float32 CUDA; corr_idx [B,N] int64 or int32 (columns of tgt);
    weight_logits [B,N]; R [B,3,3], t [B,3] = the running pose (None = identity / zero, :262-263).
    Returns (src_new [B,3,N], R_new [B,3,3], t_new [B,3], R_ab, t_ab, weights [B,N]); like the
    reference (:336-337) nothing here carries gradient."""
    _lib.require_cuda_f32(src, "src")
    _lib.require_cuda_f32(tgt, "tgt")
    _lib.require_cuda_f32(weight_logits, "weight_logits")
    if src.dim() != 3 or src.shape[1] != 3 or tgt.dim() != 3 or tgt.shape[1] != 3 or \
            tgt.shape[0] != src.shape[0]:
        raise RuntimeError("src and tgt must be [B, 3, N] and [B, 3, M], got %s and %s"
                           % (tuple(src.shape), tuple(tgt.shape)))
    B, _, N = src.shape
    M = tgt.shape[2]
    if tuple(corr_idx.shape) != (B, N) or corr_idx.dtype not in (torch.int64, torch.int32) or \
            not corr_idx.is_cuda:
        raise RuntimeError("corr_idx must be a CUDA int64/int32 tensor of shape [B, N]")
    if tuple(weight_logits.shape) != (B, N):
        raise RuntimeError("weight_logits must be [B, N]")
    src, tgt = src.detach().contiguous(), tgt.detach().contiguous()
    corr_idx, weight_logits = corr_idx.contiguous(), weight_logits.detach().contiguous()
    dev = src.device
    R_new = (torch.eye(3, device=dev).repeat(B, 1, 1) if R is None
             else R.detach().to(torch.float32).reshape(B, 3, 3).clone(memory_format=torch.contiguous_format))
    t_new = (torch.zeros(B, 3, device=dev) if t is None
             else t.detach().to(torch.float32).reshape(B, 3).clone(memory_format=torch.contiguous_format))
    src_new = torch.empty_like(src)
    R_ab = torch.empty(B, 3, 3, device=dev)
    t_ab = torch.empty(B, 3, device=dev)
    w = torch.empty(B, N, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().houv_idam_step(
            src.data_ptr(), tgt.data_ptr(), corr_idx.data_ptr(),
            1 if corr_idx.dtype == torch.int64 else 0, weight_logits.data_ptr(), B, N, M,
            src_new.data_ptr(), R_new.data_ptr(), t_new.data_ptr(), R_ab.data_ptr(),
            t_ab.data_ptr(), w.data_ptr(), _lib.stream_ptr())
    _lib.check(rc, "houv_idam_step")
    _lib.count_launches(1)
    return src_new, R_new, t_new, R_ab, t_ab, w
