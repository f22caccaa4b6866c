"""Drop-in for the functional solver of the reference's registration/train_utils.py -- the
variant behind the submission path (test.py:64, train_mult_gpu.py:108,
test_mult_modelnet.py:43):

    getPredict_angle (:359-456), solve (:467-572), rotation (:113-130), translation (:143-148),
    rotation_error / translation_error / rmse_loss (:82-95)

Same names, signatures and return values.  It is the HOUV loop with different constants, all
served by the same K3 kernel (houv_pose_optimise_ex): float64 leaf parameters and Adam state
whose float32 roundings feed the forward (:382-401), Adam lr 0.1 (:390), |T| = sin(tran_s*pi)*1
(:405), loss 6*min_1 only (:439), 500 iterations hard-coded in solve's calls (:488, :503), and
UNSEEDED NumPy draws for the initial values (:382-388; seed NumPy yourself to reproduce a run).
"""
import numpy as np
import torch

from . import solver
from .loss import rmse_loss, rotation_error, translation_error  # noqa: F401 (re-export)

ADAM_LR = 0.1            # train_utils.py:390
RETRY_THRESHOLD = 0.030  # train_utils.py:493
SOLVE_EPOCHS = 500       # train_utils.py:488,503 pass the literal 500, not solve's argument


class AverageValueMeter(object):
    """train_utils.py:22-36: running mean used by every entry script's logging."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.val = 0
        self.avg = 0
        self.sum = 0
        self.count = 0.0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


def rotation(angle, V, device='cuda'):
    """train_utils.py:113-130: Rodrigues rotation from angle[B,1] and axis V[B,3]."""
    batch_size = angle.shape[0]
    V = V / torch.sqrt((V * V).sum(dim=1, keepdim=True))
    A = torch.zeros((batch_size, 3, 3), device=V.device, dtype=V.dtype)
    A[:, 0, 1] = -V[:, 2]
    A[:, 0, 2] = V[:, 1]
    A[:, 1, 0] = V[:, 2]
    A[:, 1, 2] = -V[:, 0]
    A[:, 2, 0] = -V[:, 1]
    A[:, 2, 1] = V[:, 0]
    ones = torch.eye(3, device=V.device, dtype=V.dtype).expand(batch_size, 3, 3)
    return ones + torch.sin(angle).unsqueeze(2) * A + (1 - torch.cos(angle)).unsqueeze(2) * torch.bmm(A, A)


def translation(tran, s):
    """train_utils.py:143-148."""
    tran = tran / torch.sqrt((tran * tran).sum(dim=1, keepdim=True))
    return (tran * s).unsqueeze(1)


def initial_variables(rows):
    """The five draws of train_utils.py:382-388 from NumPy's GLOBAL generator, in the reference's
    order (V_c, angle_c, tran_c, tran_s_cpu, angle_XYZ_c -- the last one is never used but
    advances the generator) -> float64 [rows, 8] = (V_c[3], angle_c, tran_c[3], tran_s)."""
    V_c = np.random.randn(rows, 3)
    angle_c = np.random.randn(rows, 1)
    tran_c = np.random.randn(rows, 3)
    tran_s = np.random.randn(rows, 1)
    np.random.randn(rows, 3)  # angle_XYZ_c
    return np.concatenate([V_c, angle_c, tran_c, tran_s], axis=1)


def getPredict_angle(src, src_rotated, pose=None, src_ori=None, tgt_ori=None, angle_t=None,
                     label=None, kernel=64, num_epochs=1000, angle_base=0):
    """train_utils.py:359-456 -> (min_1[B,kernel], R[B,kernel,3,3], T[B,kernel,3],
    tran_s[B*kernel,1]) of the LAST forward."""
    batch_size = src.shape[0]
    rows = batch_size * kernel
    params = torch.from_numpy(initial_variables(rows)).to(src.device)
    out = solver.pose_optimise(src.float(), src_rotated.float(), params, kernel, num_epochs,
                               angle_base=angle_base, lr=ADAM_LR, wview=0.0, tran_scale=1.0,
                               tran_offset=0.0, want_tran_s=True)
    rmse_loss_ = out["min_1"].reshape((batch_size, kernel))
    R = out["R"].reshape((batch_size, kernel, 3, 3))
    T = out["T"].reshape((batch_size, kernel, 3))
    return rmse_loss_, R, T, out["tran_s"].reshape(rows, 1)


def solve(src, src_rotated, pose=None, src_ori=None, tgt_ori=None, angle_t=None, label=None,
          kernel=64, num_epochs=500, prefix='train'):
    """train_utils.py:467-572: getPredict_angle at angle_base 0, then angle_base 1..3 for the pairs
    whose best start stays above 0.030 (kept when better), best start -> ans[B,4,4] with an
    all-zero bottom row.  `num_epochs` is ignored exactly as in the reference (:488, :503)."""
    batch_size = src.shape[0]
    rmse_loss_, R, T, _ = getPredict_angle(src, src_rotated, pose, src_ori, tgt_ori, angle_t, label,
                                           kernel=kernel, num_epochs=SOLVE_EPOCHS, angle_base=0)
    rmse, _ = rmse_loss_.topk(1, dim=1, largest=False, sorted=True)
    lst_add = torch.nonzero(rmse[:, 0] > RETRY_THRESHOLD).reshape(-1)
    if int(lst_add.numel()) >= 1:
        src_add, src_rotated_add = src[lst_add], src_rotated[lst_add]
        for base_angle in range(1, 4):
            # three separate passes: each draws fresh initial values from NumPy (:382-388)
            rmse_loss_add, R_add, T_add, _ = getPredict_angle(
                src_add, src_rotated_add, None, kernel=kernel, num_epochs=SOLVE_EPOCHS,
                angle_base=base_angle)
            rmse_add, _ = rmse_loss_add.topk(1, dim=1, largest=False, sorted=True)
            flag = torch.nonzero((rmse_add < rmse[lst_add]).reshape(-1)).reshape(-1)
            geid = lst_add[flag]
            R[geid] = R_add[flag]
            rmse_loss_[geid] = rmse_loss_add[flag]
            T[geid] = T_add[flag]
            rmse[geid] = rmse_add[flag]
    best = rmse_loss_.argmin(dim=1)
    rows = torch.arange(batch_size, device=src.device)
    ans = torch.zeros((batch_size, 4, 4), device=src.device)
    ans[:, :3, :3] = R[rows, best]
    ans[:, :3, 3] = T[rows, best]
    if prefix == 'test':
        return ans.cpu()
    r_err = rotation_error(ans[:, :3, :3], pose[:, :3, :3])
    t_err = translation_error(ans[:, :3, 3], pose[:, :3, 3])
    print(r_err.mean(), t_err.mean(), rmse.max())
    return r_err, t_err, ans
