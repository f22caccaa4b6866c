"""Loss helpers and metrics of the HOUV path, same names and semantics as the
reference (paths relative to /root/reference):
    calc_cd, calc_cd_percent, calc_cd_percent_aligned, loss_view
                         registration/model_utils_completion.py:69-116,157-166
                         (calc_cd also completion/model_utils.py:69-80)
    rotation_error, translation_error, rmse_loss     registration/train_utils.py:82-95
    Predict_loss                                      registration/models/houv.py:209-222
These are the UNFUSED building blocks (torch ops around the K1/K2 operator) for
callers that assemble their own loss; predict_model uses the fused K3 kernel
and is tested to agree with them.
"""
import math

import torch

from .chamfer import cd, fscore
from .emd import calc_emd, emd  # noqa: F401  (model_utils_completion.py:170-175 lives beside calc_cd)


def calc_cd(output, gt, calc_f1=False):
    if not (torch.is_grad_enabled() and (output.requires_grad or gt.requires_grad)):
        return calc_cd_fused(output, gt, calc_f1)  # evaluation: forward + one reduction launch
    cham_loss = cd()
    dist1, dist2, _, _ = cham_loss(gt, output)
    cd_p = (torch.sqrt(dist1).mean(1) + torch.sqrt(dist2).mean(1)) / 2
    cd_t = (dist1.mean(1) + dist2.mean(1))
    if calc_f1:
        f1, _, _ = fscore(dist1, dist2)
        return cd_p, cd_t, f1
    return cd_p, cd_t


def calc_cd_fused(output, gt, calc_f1=False, threshold=0.0001):
    """calc_cd without autograd (the evaluation calls, e.g. completion/models/vrcnet.py:647):
    Chamfer forward (K1) + houv_cd_reduce, two launches in all."""
    from . import _lib
    dist1, dist2, _, _ = cd()(gt.detach(), output.detach())
    B, N = dist1.shape
    M = dist2.shape[1]
    out = torch.empty(3, B, device=dist1.device)
    with torch.cuda.device(dist1.device):
        rc = _lib.load().houv_cd_reduce(dist1.data_ptr(), dist2.data_ptr(), B, N, M, threshold,
                                        out[0].data_ptr(), out[1].data_ptr(),
                                        out[2].data_ptr() if calc_f1 else 0, _lib.stream_ptr())
    _lib.check(rc, "houv_cd_reduce")
    _lib.count_launches(1)
    return (out[0], out[1], out[2]) if calc_f1 else (out[0], out[1])


def calc_cd_percent(output, gt, calc_f1=False, percent=1):
    point = output.shape[1]
    k = int(point * percent)
    cham_loss = cd()
    dist1, dist2, _, _ = cham_loss(gt, output)
    dist1, _ = dist1.topk(k, dim=1, largest=False, sorted=True)
    dist2, _ = dist2.topk(k, dim=1, largest=False, sorted=True)
    cd_p = torch.sqrt(dist1).mean(1)
    cd_t = torch.sqrt(dist2).mean(1)
    if calc_f1:
        f1, _, _ = fscore(dist1, dist2)
        return cd_p, cd_t, f1
    return cd_p, cd_t


def calc_cd_percent_aligned(output, gt, percent=1):
    point = output.shape[1]
    k = int(point * percent)
    cham_loss = cd()
    dist1, dist2, idx1, idx2 = cham_loss(gt, output)
    dist1, idxx1 = dist1.topk(k, dim=1, largest=False, sorted=True)
    dist2, idxx2 = dist2.topk(k, dim=1, largest=False, sorted=True)
    cd_p = torch.sqrt(dist1).mean(1)
    cd_t = torch.sqrt(dist2).mean(1)
    return cd_p, cd_t, idx1, idx2, idxx1, idxx2


def loss_view(src, tgt, dim=0, percent=1):
    ones = torch.ones((1, 1, 3), device=src.device, dtype=src.dtype)
    ones[:, :, dim] = 0
    mask = torch.zeros_like(src) + ones
    a = src * mask
    tgt = tgt * mask
    cd_t, cd_p = calc_cd_percent(a, tgt, percent=percent)
    return cd_t, cd_p


def Predict_loss(src, src_rotated, alpha=0.5):
    cd_t, cd_p = calc_cd_percent(src, src_rotated, percent=alpha)
    min_1, _ = torch.min(torch.cat([cd_t.unsqueeze(1), cd_p.unsqueeze(1)], dim=1), dim=1)
    cd_t_view_0, cd_p_view_0 = loss_view(src, src_rotated, dim=0)
    cd_t_view_1, cd_p_view_1 = loss_view(src, src_rotated, dim=1)
    cd_t_view_2, cd_p_view_2 = loss_view(src, src_rotated, dim=2)
    min_v_0, _ = torch.min(torch.cat([cd_t_view_0.unsqueeze(1), cd_p_view_0.unsqueeze(1)], dim=1), dim=1)
    min_v_1, _ = torch.min(torch.cat([cd_t_view_1.unsqueeze(1), cd_p_view_1.unsqueeze(1)], dim=1), dim=1)
    min_v_2, _ = torch.min(torch.cat([cd_t_view_2.unsqueeze(1), cd_p_view_2.unsqueeze(1)], dim=1), dim=1)
    return (min_1) * 6 + (min_v_0 + min_v_1 + min_v_2), min_1


def rotation_error(R, R_gt):
    cos_theta = (torch.einsum('bij,bij->b', R, R_gt) - 1) / 2
    cos_theta = torch.clamp(cos_theta, -1, 1)
    return torch.acos(cos_theta) * 180 / math.pi


def translation_error(t, t_gt):
    return torch.norm(t - t_gt, dim=1)


def rmse_loss(pts, T, T_gt):
    pts_pred = pts @ T[:, :3, :3].transpose(1, 2) + T[:, :3, 3].unsqueeze(1)
    pts_gt = pts @ T_gt[:, :3, :3].transpose(1, 2) + T_gt[:, :3, 3].unsqueeze(1)
    return torch.norm(pts_pred - pts_gt, dim=2).mean(dim=1)
