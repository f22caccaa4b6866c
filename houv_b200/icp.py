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
