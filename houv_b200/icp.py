"""ICP baseline of the reference (registration/train_ICP.py:124-190) on the GPU.

The reference calls Open3D 0.9.0 per pair on the host:
    o3d.registration.registration_icp(pcd, pcd2, threshold=0.02, trans_init,
        TransformationEstimationPointToPoint(), ICPConvergenceCriteria(max_iteration=500))
with the fixed initial rotation of :138-140.  `registration_icp` below keeps that
call shape for a whole batch (one CTA per pair, houv_icp); `icp_batch` is the
body of train_ICP.train()'s loop: ICP for every pair, then the metrics of :187-190.
"""
import numpy as np
import torch

from . import _lib
from .loss import rmse_loss, rotation_error, translation_error

# train_ICP.py:138-140
TRANS_INIT = np.asarray([[0.862, 0.011, -0.507, 0.0],
                         [-0.139, 0.967, -0.215, 0.0],
                         [0.487, 0.255, 0.835, 0.0],
                         [0.0, 0.0, 0.0, 1.0]], dtype=np.float32)
THRESHOLD = 0.02        # train_ICP.py:124
MAX_ITERATION = 500     # train_ICP.py:151


class RegistrationResult:
    """Batched stand-in for o3d.registration.RegistrationResult."""

    def __init__(self, transformation, fitness, inlier_rmse, iterations, correspondences):
        self.transformation = transformation
        self.fitness = fitness
        self.inlier_rmse = inlier_rmse
        self.iterations = iterations
        self.correspondences = correspondences


RESIDENT_MAX_POINTS = 2048  # csrc/icp.cu kIcpPts: one CTA keeps a pair resident for all iterations


def _registration_icp_large(source, target, max_distance, init, max_iteration, relative_fitness,
                            relative_rmse):
    """The same loop for clouds above RESIDENT_MAX_POINTS, which the resident kernel cannot hold:
    per iteration one directed search of K1 (any size) over the pairs still iterating, the gated
    sums and the Kabsch update in float64 on the device (batched SVD), convergence per pair.
    Open3D takes any cloud size (train_ICP.py:148-151), so does this call."""
    from .chamfer import nn_search
    P, N, _ = source.shape
    dev = source.device
    T = (init.double() if init.dim() == 3 else init.double().expand(P, 4, 4)).clone()
    thr2 = float(max_distance) ** 2
    srcd = source.double()

    def evaluate(idx, Tm):
        y = torch.bmm(srcd[idx], Tm[:, :3, :3].transpose(1, 2)) + Tm[:, :3, 3].unsqueeze(1)
        d, j = nn_search(y.float().contiguous(), target[idx].contiguous())
        keep = (d < thr2)
        n = keep.sum(1)
        fit = n.double() / N
        rm = torch.sqrt((d.double() * keep).sum(1) / n.clamp(min=1).double()) * (n > 0)
        w = torch.gather(target[idx].double(), 1, j.long().unsqueeze(2).expand(-1, -1, 3))
        return fit, rm, n, y, w, keep.double()

    every = torch.arange(P, device=dev)
    fit, rm, ncorr, y, w, keep = evaluate(every, T)
    its = torch.zeros(P, device=dev)
    idx = every
    for i in range(max_iteration):
        if idx.numel() == 0:
            break
        n = keep.sum(1, keepdim=True).clamp(min=1.0)
        ym, wm = (y * keep.unsqueeze(2)).sum(1) / n, (w * keep.unsqueeze(2)).sum(1) / n
        H = torch.bmm(((y - ym.unsqueeze(1)) * keep.unsqueeze(2)).transpose(1, 2), w - wm.unsqueeze(1))
        U, _, Vt = torch.linalg.svd(H)
        V = Vt.transpose(1, 2)
        D = torch.eye(3, device=dev, dtype=torch.float64).repeat(idx.numel(), 1, 1)
        D[:, 2, 2] = torch.where(torch.det(torch.bmm(V, U.transpose(1, 2))) < 0, -1.0, 1.0)
        R = torch.bmm(torch.bmm(V, D), U.transpose(1, 2))
        upd = torch.eye(4, device=dev, dtype=torch.float64).repeat(idx.numel(), 1, 1)
        upd[:, :3, :3] = R
        upd[:, :3, 3] = wm - torch.bmm(R, ym.unsqueeze(2)).squeeze(2)
        has = keep.sum(1) > 0          # no correspondence: the transformation stays (Open3D returns early)
        Tn = torch.where(has.reshape(-1, 1, 1), torch.bmm(upd, T[idx]), T[idx])
        T[idx] = Tn
        nfit, nrm, nn_, y, w, keep = evaluate(idx, Tn)
        stop = ((fit[idx] - nfit).abs() < relative_fitness) & ((rm[idx] - nrm).abs() < relative_rmse)
        fit[idx], rm[idx], ncorr[idx], its[idx] = nfit, nrm, nn_, float(i + 1)
        go = ~stop
        idx, y, w, keep = idx[go], y[go], w[go], keep[go]
    return RegistrationResult(T.float(), fit.float(), rm.float(), its, ncorr.float())


def registration_icp(source, target, max_correspondence_distance=THRESHOLD, init=None,
                     max_iteration=MAX_ITERATION, relative_fitness=1e-6, relative_rmse=1e-6):
    """source[P,N,3], target[P,M,3] float32 CUDA; init: 4x4 (shared) or [P,4,4]."""
    _lib.require_cuda_f32(source, "source", 3)
    _lib.require_cuda_f32(target, "target", 3)
    source, target = source.contiguous(), target.contiguous()
    P, N, _ = source.shape
    M = target.shape[1]
    if init is None:
        init = np.eye(4, dtype=np.float32)
    init = torch.as_tensor(init, dtype=torch.float32).to(source.device).contiguous()
    per_pair = 1 if init.dim() == 3 else 0
    if per_pair and init.shape[0] != P:
        raise RuntimeError("init must be [4,4] or [P,4,4]")
    if max(N, M) > RESIDENT_MAX_POINTS and P > 0:
        return _registration_icp_large(source, target, max_correspondence_distance, init,
                                       max_iteration, relative_fitness, relative_rmse)
    T = torch.empty(P, 4, 4, device=source.device)
    stats = torch.empty(P, 4, device=source.device)
    with torch.cuda.device(source.device):
        rc = _lib.load().houv_icp(source.data_ptr(), target.data_ptr(), P, N, M, init.data_ptr(),
                                  per_pair, float(max_correspondence_distance), int(max_iteration),
                                  float(relative_fitness), float(relative_rmse), T.data_ptr(),
                                  stats.data_ptr(), _lib.stream_ptr())
    _lib.check(rc, "houv_icp")
    _lib.count_launches(1)
    return RegistrationResult(T, stats[:, 0], stats[:, 1], stats[:, 2], stats[:, 3])


def icp_batch(src_rotated, tgt_rotated, transform=None, threshold=THRESHOLD, trans_init=TRANS_INIT,
              max_iteration=MAX_ITERATION):
    """train_ICP.py:137-190 for one batch: ans[B,4,4] and (r_err, t_err, mse) when the ground
    truth `transform` is given."""
    reg = registration_icp(src_rotated, tgt_rotated, threshold, trans_init, max_iteration)
    ans = reg.transformation
    if transform is None:
        return ans
    r_err = rotation_error(ans[:, :3, :3], transform[:, :3, :3])
    t_err = translation_error(ans[:, :3, 3], transform[:, :3, 3])
    mse = rmse_loss(src_rotated, ans, transform)
    return ans, r_err, t_err, mse
