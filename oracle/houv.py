"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by houv_b200/).

CPU restatement (torch-CPU autograd + the C chamfer oracle) of the reference's
HOUV solver:
    registration/models/houv.py:13-222           HOUV, predict_model, solve_model, Predict_loss
    registration/model_utils_completion.py:69-166 calc_cd, calc_cd_percent, loss_view
    registration/train_utils.py:82-95             rotation_error, translation_error, rmse_loss
    registration/train_utils.py:113-148,359-572   rotation, translation, getPredict_angle, solve
    registration/model_utils.py:213-255           SVDHead
(paths relative to /root/reference).  The reference files hard-code .cuda() and
import absent packages, so they cannot run on the GPU box; this restatement can.

Parity status: PINNED for everything above -- tests/golden/make_golden.py runs
the reference's own python files, unmodified, in the dev container
(tests/golden/ref_loader.py) and tests/test_oracle_cpu.py checks this module
against those outputs bit for bit (same torch ops in the same order).
UNPINNED: icp_point_to_point (Open3D 0.9.0 is third-party and absent; see its
docstring).
"""
import math

import numpy as np
import torch
from torch import nn

from . import chamfer as oc


# ----------------------------------------------------------------- chamfer op
class ChamferFunction(torch.autograd.Function):
    """dist_chamfer_3D.py:26-64 on the CPU through oracle/chamfer_oracle.c."""

    @staticmethod
    def forward(ctx, xyz1, xyz2):
        d1, d2, i1, i2 = oc.chamfer_forward(xyz1.detach().numpy(), xyz2.detach().numpy())
        idx1, idx2 = torch.from_numpy(i1), torch.from_numpy(i2)
        ctx.save_for_backward(xyz1, xyz2, idx1, idx2)
        return torch.from_numpy(d1), torch.from_numpy(d2), idx1, idx2

    @staticmethod
    def backward(ctx, graddist1, graddist2, gradidx1, gradidx2):
        xyz1, xyz2, idx1, idx2 = ctx.saved_tensors
        g1, g2 = oc.chamfer_backward(xyz1.detach().numpy(), xyz2.detach().numpy(),
                                     graddist1.contiguous().numpy(),
                                     graddist2.contiguous().numpy(), idx1.numpy(), idx2.numpy())
        return torch.from_numpy(g1), torch.from_numpy(g2)


def cd(input1, input2):
    """chamfer_3DDist().forward, dist_chamfer_3D.py:71-74."""
    return ChamferFunction.apply(input1.contiguous().float(), input2.contiguous().float())


# --------------------------------------------------------------- loss helpers
def fscore(dist1, dist2, threshold=0.0001):
    """utils/metrics/CD/fscore.py:3-16."""
    precision_1 = torch.mean((dist1 < threshold).float(), dim=1)
    precision_2 = torch.mean((dist2 < threshold).float(), dim=1)
    f = 2 * precision_1 * precision_2 / (precision_1 + precision_2)
    f[torch.isnan(f)] = 0
    return f, precision_1, precision_2


def calc_cd(output, gt, calc_f1=False):
    """model_utils_completion.py:69-80 = completion/model_utils.py:69-80."""
    dist1, dist2, _, _ = cd(gt, output)
    cd_p = (torch.sqrt(dist1).mean(1) + torch.sqrt(dist2).mean(1)) / 2
    cd_t = dist1.mean(1) + dist2.mean(1)
    if calc_f1:
        f1, _, _ = fscore(dist1, dist2)
        return cd_p, cd_t, f1
    return cd_p, cd_t


def calc_cd_percent(output, gt, percent=1):
    """model_utils_completion.py:83-100. NB cham_loss(gt, output): xyz1 = gt."""
    point = output.shape[1]
    k = int(point * percent)
    dist1, dist2, _, _ = cd(gt, output)
    dist1, _ = dist1.topk(k, dim=1, largest=False, sorted=True)
    dist2, _ = dist2.topk(k, dim=1, largest=False, sorted=True)
    cd_p = torch.sqrt(dist1).mean(1)
    cd_t = torch.sqrt(dist2).mean(1)
    return cd_p, cd_t


def loss_view(src, tgt, dim=0, percent=1):
    """model_utils_completion.py:157-166."""
    mask = torch.zeros_like(src)
    ones = torch.ones((1, 1, 3))
    ones[:, :, dim] = 0
    mask = mask + ones
    a = src * mask
    tgt = tgt * mask
    cd_t, cd_p = calc_cd_percent(a, tgt, percent=percent)
    return cd_t, cd_p


def Predict_loss(src, src_rotated, alpha=0.5):
    """houv.py:209-222."""
    cd_t, cd_p = calc_cd_percent(src, src_rotated, percent=alpha)
    min_1, _ = torch.min(torch.cat([cd_t.unsqueeze(1), cd_p.unsqueeze(1)], dim=1), dim=1)
    cd_t_view_0, cd_p_view_0 = loss_view(src, src_rotated, dim=0)
    cd_t_view_1, cd_p_view_1 = loss_view(src, src_rotated, dim=1)
    cd_t_view_2, cd_p_view_2 = loss_view(src, src_rotated, dim=2)
    min_v_0, _ = torch.min(torch.cat([cd_t_view_0.unsqueeze(1), cd_p_view_0.unsqueeze(1)], dim=1), dim=1)
    min_v_1, _ = torch.min(torch.cat([cd_t_view_1.unsqueeze(1), cd_p_view_1.unsqueeze(1)], dim=1), dim=1)
    min_v_2, _ = torch.min(torch.cat([cd_t_view_2.unsqueeze(1), cd_p_view_2.unsqueeze(1)], dim=1), dim=1)
    return (min_1) * 6 + (min_v_0 + min_v_1 + min_v_2), min_1


# -------------------------------------------------------------------- metrics
def rotation_error(R, R_gt):
    """train_utils.py:82-85 (degrees)."""
    cos_theta = (torch.einsum('bij,bij->b', R, R_gt) - 1) / 2
    cos_theta = torch.clamp(cos_theta, -1, 1)
    return torch.acos(cos_theta) * 180 / math.pi


def translation_error(t, t_gt):
    """train_utils.py:88-89."""
    return torch.norm(t - t_gt, dim=1)


def rmse_loss(pts, T, T_gt):
    """train_utils.py:92-95."""
    pts_pred = pts @ T[:, :3, :3].transpose(1, 2) + T[:, :3, 3].unsqueeze(1)
    pts_gt = pts @ T_gt[:, :3, :3].transpose(1, 2) + T_gt[:, :3, 3].unsqueeze(1)
    return torch.norm(pts_pred - pts_gt, dim=2).mean(dim=1)


# ------------------------------------------------------------ model + solver
def initial_parameters(rows, seed=2021):
    """HOUV.reset_weight's draws, houv.py:40-61: numpy is re-seeded before EVERY
    tensor, so angle_c == tran_s and tran_c reuses the V_c stream; the first 26
    GLOBAL rows of V_c are the axis set {-1,0,1}^3 \\ 0.  -> float32 arrays
    V_c[rows,3], angle_c[rows,1], tran_c[rows,3], tran_s[rows,1]."""
    np.random.seed(seed)
    vc = np.random.randn(rows, 3)
    num = 0
    for x0 in [-1, 0, 1]:
        for y0 in [-1, 0, 1]:
            for z0 in [-1, 0, 1]:
                if x0 == 0 and y0 == 0 and z0 == 0:
                    continue
                vc[num] = np.array((x0, y0, z0))  # IndexError if rows < 26, like the reference
                num += 1
    np.random.seed(seed)
    np.random.seed(seed)
    angle_c = np.random.randn(rows, 1)
    np.random.seed(seed)
    tran_c = np.random.randn(rows, 3)
    np.random.seed(seed)
    tran_s = np.random.randn(rows, 1)
    f = np.float32
    return vc.astype(f), angle_c.astype(f), tran_c.astype(f), tran_s.astype(f)


class HOUV(nn.Module):
    """houv.py:13-103 without the .cuda() calls (and without the stale
    tran_s_cpu parameter of __init__, :36, which never receives a gradient)."""

    def __init__(self, batch_size=0, angle_base=0):
        super().__init__()
        self.batch_size = batch_size
        self.angle_base = angle_base
        self.pi = torch.acos(torch.zeros(1)).item() * 2  # float32 pi, houv.py:19

    def reset_weight(self, batch_size, angle_base, seed=2021):
        self.batch_size = batch_size
        vc, ac, tc, ts = initial_parameters(batch_size, seed)
        self.angle_base = angle_base
        self.V_c = nn.Parameter(torch.from_numpy(vc), requires_grad=True)
        self.angle_c = nn.Parameter(torch.from_numpy(ac), requires_grad=True)
        self.tran_c = nn.Parameter(torch.from_numpy(tc), requires_grad=True)
        self.tran_s = nn.Parameter(torch.from_numpy(ts), requires_grad=True)
        self.ones = torch.zeros((batch_size, 3, 3)) + torch.eye(3)

    def cd_rotation(self, angle, V):
        """houv.py:69-86 (Rodrigues from the normalised axis)."""
        batch_size = angle.shape[0]
        V = V / torch.sqrt((V * V).sum(dim=1, keepdim=True))
        A = torch.zeros((batch_size, 3, 3), dtype=torch.float)
        A[:, 0, 1] = -V[:, 2]
        A[:, 0, 2] = V[:, 1]
        A[:, 1, 0] = V[:, 2]
        A[:, 1, 2] = -V[:, 0]
        A[:, 2, 0] = -V[:, 1]
        A[:, 2, 1] = V[:, 0]
        R = self.ones + torch.sin(angle).unsqueeze(2) * A + \
            (1 - torch.cos(angle)).unsqueeze(2) * torch.bmm(A, A)
        return R

    def translation(self, tran, s):
        """houv.py:88-92."""
        tran = tran / torch.sqrt((tran * tran).sum(dim=1, keepdim=True))
        tran = tran * s
        return tran.unsqueeze(1)

    def forward(self, src):
        """houv.py:94-103."""
        src = src.squeeze(0)
        angle = torch.sin(self.angle_c * self.pi) * self.pi / 8 + self.pi / 8 + \
            self.angle_base * self.pi / 4
        R = self.cd_rotation(angle, self.V_c)
        tran_s = torch.sin(self.tran_s * self.pi) * 0.125 + 0.125
        T = self.translation(self.tran_c, tran_s)
        src_temp = torch.bmm(src, R.transpose(1, 2)) + T
        return src_temp, R, T


def predict_model(net, src, src_rotated, kernel=64, num_epochs=500, angle_base=0, seed=2021,
                  trace=None):
    """houv.py:106-138.  `trace` (a list) receives per-iteration
    (loss_per_row, min_1, params) copies for trajectory tests."""
    batch_size = src.shape[0]
    num_points = src.shape[1]
    src = src.unsqueeze(1).expand((-1, kernel, -1, -1)).reshape(-1, num_points, 3)
    src_rotated = src_rotated.unsqueeze(1).expand((-1, kernel, -1, -1)).reshape(
        -1, src_rotated.shape[1], 3)
    batch_size *= kernel
    net.reset_weight(batch_size, angle_base, seed=seed)
    optimizer = torch.optim.Adam(net.parameters(), lr=0.01)
    net.train()
    for i in range(num_epochs):
        optimizer.zero_grad()
        src_temp, R, T = net(src)
        rmse_loss_, min_1 = Predict_loss(src_temp, src_rotated)
        rmse = rmse_loss_.mean()
        rmse.backward()
        if trace is not None:
            trace.append(dict(
                loss=rmse_loss_.detach().clone(), min_1=min_1.detach().clone(),
                grads=[p.grad.detach().clone() for p in (net.V_c, net.angle_c, net.tran_c, net.tran_s)],
                params=[p.detach().clone() for p in (net.V_c, net.angle_c, net.tran_c, net.tran_s)]))
        optimizer.step()
    batch_size = batch_size // kernel
    rmse_loss_ = (min_1).reshape((batch_size, kernel))
    R = R.reshape((batch_size, kernel, 3, 3))
    T = T.reshape((batch_size, kernel, 3))
    return rmse_loss_, R, T


def solve_model(net, src, src_rotated, pose=None, kernel=64, num_epochs=200, prefix='train',
                threshold=0.030):
    """houv.py:142-206: angle_base 0 for everyone, retries at 1..3 for pairs whose
    best min_1 stays above 0.030, best start -> ans[B,4,4] with a ZERO bottom row."""
    batch_size = src.shape[0]
    rmse_loss_, R, T = predict_model(net, src, src_rotated, kernel=kernel,
                                     num_epochs=num_epochs, angle_base=0)
    rmse_loss_, R, T = rmse_loss_.detach().clone(), R.detach().clone(), T.detach().clone()
    rmse, b_index = rmse_loss_.topk(1, dim=1, largest=False, sorted=True)
    lst_add = [j for j in range(batch_size) if rmse[j][0] > threshold]
    if len(lst_add) > 0:
        lst_add = np.array(lst_add).astype(int)
        src_add, src_rotated_add = src[lst_add], src_rotated[lst_add]
        for base_angle in range(1, 4):
            la, Ra, Ta = predict_model(net, src_add, src_rotated_add, kernel=kernel,
                                       num_epochs=num_epochs, angle_base=base_angle)
            la, Ra, Ta = la.detach(), Ra.detach(), Ta.detach()
            rmse_add, _ = la.topk(1, dim=1, largest=False, sorted=True)
            flag = (rmse_add < rmse[lst_add]).reshape(-1)
            flag = torch.nonzero(flag).reshape(-1)
            geid = lst_add[flag.int().numpy()]
            R[geid] = Ra[flag]
            rmse_loss_[geid] = la[flag]
            T[geid] = Ta[flag]
            rmse[geid] = rmse_add[flag]
    ans = torch.zeros((batch_size, 4, 4))
    for i in range(batch_size):
        rm, id_ = rmse_loss_[i].topk(1, dim=0, largest=False, sorted=True)
        ans[i, :3, :3] = R[i][id_]
        ans[i, :3, 3] = T[i][id_]
    if prefix == 'test':
        return ans
    r_err = rotation_error(ans[:, :3, :3], pose[:, :3, :3])
    t_err = translation_error(ans[:, :3, 3], pose[:, :3, 3])
    return r_err, t_err, ans


# ---------------------------------------------------------------- Kabsch head
# ------------------------------------------------- train_utils.py functional solver
def tu_rotation(angle, V):
    """train_utils.py:113-130."""
    batch_size = angle.shape[0]
    V = V / torch.sqrt((V * V).sum(dim=1, keepdim=True))
    A = torch.zeros((batch_size, 3, 3))
    A[:, 0, 1] = -V[:, 2]
    A[:, 0, 2] = V[:, 1]
    A[:, 1, 0] = V[:, 2]
    A[:, 1, 2] = -V[:, 0]
    A[:, 2, 0] = -V[:, 1]
    A[:, 2, 1] = V[:, 0]
    ones = torch.zeros((batch_size, 3, 3)) + torch.eye(3)
    return ones + torch.sin(angle).unsqueeze(2) * A + (1 - torch.cos(angle)).unsqueeze(2) * torch.bmm(A, A)


def tu_translation(tran, s):
    """train_utils.py:143-148."""
    tran = tran / torch.sqrt((tran * tran).sum(dim=1, keepdim=True))
    return (tran * s).unsqueeze(1)


def getPredict_angle(src, src_rotated, kernel=64, num_epochs=1000, angle_base=0, trace=None):
    """train_utils.py:359-456: float64 leaves drawn from NumPy's global generator, their float32
    roundings in the forward, Adam(lr=0.1), loss 6*min_1, |T| = sin(tran_s*pi)*1.
    -> (min_1[B,K], R[B,K,3,3], T[B,K,3], tran_s[B*K,1]) of the last forward.
    trace (list) receives per iteration dict(loss, min_1, grads[4])."""
    batch_size, num_points = src.shape[0], src.shape[1]
    src = src.unsqueeze(1).expand((-1, kernel, -1, -1)).reshape(-1, num_points, 3)
    src_rotated = src_rotated.unsqueeze(1).expand((-1, kernel, -1, -1)).reshape(-1, num_points, 3)
    batch_size *= kernel
    V_c = torch.from_numpy(np.random.randn(batch_size, 3)).requires_grad_()
    angle_c = torch.from_numpy(np.random.randn(batch_size, 1)).requires_grad_()
    tran_c = torch.from_numpy(np.random.randn(batch_size, 3)).requires_grad_()
    tran_s_cpu = torch.from_numpy(np.random.randn(batch_size, 1)).requires_grad_()
    angle_XYZ_c = torch.from_numpy(np.random.randn(batch_size, 3)).requires_grad_()
    vari_list = [V_c, angle_c, tran_c, tran_s_cpu, angle_XYZ_c]
    optimizer = torch.optim.Adam(vari_list, lr=0.1)
    pi = torch.acos(torch.zeros(1)).item() * 2
    for i in range(num_epochs):
        optimizer.zero_grad()
        V, tran, angle, tran_s = V_c.float(), tran_c.float(), angle_c.float(), tran_s_cpu.float()
        angle = torch.sin(angle * pi) * pi / 8 + pi / 8 + angle_base * pi / 4
        tran_s = torch.sin(tran_s * pi) * 1
        R = tu_rotation(angle, V)
        T = tu_translation(tran, tran_s)
        src_temp = torch.bmm(src, R.transpose(1, 2)) + T
        cd_t, cd_p = calc_cd_percent(src_temp, src_rotated, percent=0.5)
        min_1, _ = torch.min(torch.cat([cd_t.unsqueeze(1), cd_p.unsqueeze(1)], dim=1), dim=1)
        rmse_loss_ = min_1 * 6
        rmse_loss_.mean().backward()
        if trace is not None:
            trace.append(dict(loss=rmse_loss_.detach().clone(), min_1=min_1.detach().clone(),
                              grads=[p.grad.detach().clone() for p in vari_list[:4]]))
        optimizer.step()
    batch_size = batch_size // kernel
    return (min_1.reshape((batch_size, kernel)), R.reshape((batch_size, kernel, 3, 3)),
            T.reshape((batch_size, kernel, 3)), tran_s)


def solve(src, src_rotated, pose=None, kernel=64, epochs=500, prefix='train', threshold=0.030):
    """train_utils.py:467-572 (angle_num = 1).  `epochs` stands for the literal 500 the reference
    passes to getPredict_angle (:488, :503)."""
    batch_size = src.shape[0]
    rmse_loss_, R, T, _ = getPredict_angle(src, src_rotated, kernel=kernel, num_epochs=epochs,
                                           angle_base=0)
    rmse_loss_, R, T = rmse_loss_.detach().clone(), R.detach().clone(), T.detach().clone()
    rmse, _ = rmse_loss_.topk(1, dim=1, largest=False, sorted=True)
    lst_add = [j for j in range(batch_size) if rmse[j][0] > threshold]
    if len(lst_add) >= 1:
        src_add, src_rotated_add = src[lst_add], src_rotated[lst_add]
        lst_add = np.array(lst_add).astype(int)
        for base_angle in range(1, 4):
            rmse_loss_add, R_add, T_add, _ = getPredict_angle(src_add, src_rotated_add, kernel=kernel,
                                                              num_epochs=epochs, angle_base=base_angle)
            rmse_add, _ = rmse_loss_add.detach().topk(1, dim=1, largest=False, sorted=True)
            flag = torch.nonzero((rmse_add < rmse[lst_add]).reshape(-1)).reshape(-1)
            geid = lst_add[flag.int().numpy()]
            R[geid] = R_add.detach()[flag]
            rmse_loss_[geid] = rmse_loss_add.detach()[flag]
            T[geid] = T_add.detach()[flag]
            rmse[geid] = rmse_add[flag]
    ans = torch.zeros((batch_size, 4, 4))
    for i in range(batch_size):
        _, idx = rmse_loss_[i].topk(1, dim=0, largest=False, sorted=True)
        ans[i, :3, :3] = R[i][idx]
        ans[i, :3, 3] = T[i][idx]
    if prefix == 'test':
        return ans
    return (rotation_error(ans[:, :3, :3], pose[:, :3, :3]),
            translation_error(ans[:, :3, 3], pose[:, :3, 3]), ans)


class SVDHead(nn.Module):
    """registration/model_utils.py:213-255, restated verbatim in behaviour (per-sample
    torch.svd loop, reflection fix, weighted translation)."""

    def __init__(self):
        super().__init__()
        self.reflect = nn.Parameter(torch.eye(3), requires_grad=False)
        self.reflect[2, 2] = -1

    def forward(self, src, src_corr, weights=None):
        src_centered = src - src.mean(dim=2, keepdim=True)
        src_corr_centered = src_corr - src_corr.mean(dim=2, keepdim=True)
        if weights is None:
            H = torch.matmul(src_centered, src_corr_centered.transpose(2, 1).contiguous())
        else:
            H = torch.matmul(src_centered * weights, src_corr_centered.transpose(2, 1).contiguous())
        R = []
        for i in range(src.size(0)):
            u, s, v = torch.svd(H[i])
            r = torch.matmul(v, u.transpose(1, 0).contiguous())
            if torch.det(r) < 0:
                u, s, v = torch.svd(H[i])
                v = torch.matmul(v, self.reflect)
                r = torch.matmul(v, u.transpose(1, 0).contiguous())
            R.append(r)
        R = torch.stack(R, dim=0)
        if weights is None:
            t = torch.matmul(-R, src.mean(dim=2, keepdim=True)) + src_corr.mean(dim=2, keepdim=True)
        else:
            t = torch.matmul(-R, (weights * src).sum(dim=2, keepdim=True)) + \
                (weights * src_corr).sum(dim=2, keepdim=True)
        return R, t.view(src.size(0), 3)


def idam_step(src, tgt, corr_idx, weight_logits, R, t, head=None):
    """The tail of one iteration of IDAM.forward (registration/models/idam.py:317-342), torch CPU:
    correspondences by index (:318), hybrid point elimination (:329-331), weighted SVDHead (:335),
    move src and compose the running pose (:338-340).  `head` = the SVD head to use (the reference's
    own class when making golden vectors, ours above otherwise)."""
    head = head or SVDHead()
    src_corr = tgt[np.arange(tgt.size(0))[:, np.newaxis], :, corr_idx].transpose(1, 2)
    weights = torch.sigmoid(weight_logits)
    weights = weights * (weights >= weights.median(-1, keepdim=True)[0]).float()
    weights = weights / (weights.sum(-1, keepdim=True) + 1e-8)
    rotation_ab, translation_ab = head(src, src_corr, weights.unsqueeze(1))
    src = torch.matmul(rotation_ab, src) + translation_ab.unsqueeze(-1)
    R = torch.matmul(rotation_ab, R)
    t = torch.matmul(rotation_ab, t.unsqueeze(-1)).squeeze(-1) + translation_ab
    return src, R, t, rotation_ab, translation_ab, weights


# ------------------------------------------------------------------------ ICP
def icp_point_to_point(source, target, max_distance=0.02, init=None, max_iteration=500,
                       relative_fitness=1e-6, relative_rmse=1e-6):
    """PARITY UNPINNED.  float64 restatement of Open3D 0.9.0's
    registration_icp(..., TransformationEstimationPointToPoint(),
    ICPConvergenceCriteria(max_iteration)) as called at registration/train_ICP.py:148-151.
    Open3D (pinned at 0.9.0 by completion/requirements.txt:4) is neither in the reference tree
    nor installed, and no reference test or fixture exercises it, so this follows the
    published algorithm:
        result = evaluate(init)
        for i < max_iteration: update = umeyama_no_scale(correspondences); T = update @ T;
                               result' = evaluate(T); stop when both |d fitness| and
                               |d inlier_rmse| fall below the relative criteria
        evaluate: nearest target of each transformed source point, a correspondence when the
                  squared distance < max_distance**2; fitness = #corr/#source,
                  inlier_rmse = sqrt(sum d^2/#corr)
    source[N,3], target[M,3] numpy -> (T[4,4] float64, fitness, inlier_rmse, iterations)."""
    src = np.asarray(source, np.float64)
    tgt = np.asarray(target, np.float64)
    T = np.eye(4) if init is None else np.asarray(init, np.float64).copy()

    def evaluate(T):
        y = src @ T[:3, :3].T + T[:3, 3]
        d2 = ((y[:, None, :] - tgt[None, :, :]) ** 2).sum(-1)
        j = d2.argmin(1)
        d = d2[np.arange(len(y)), j]
        keep = d < max_distance ** 2
        n = int(keep.sum())
        fit = n / float(len(y))
        rm = float(np.sqrt(d[keep].sum() / n)) if n else 0.0
        return fit, rm, y[keep], tgt[j[keep]]

    fit, rm, ys, ws = evaluate(T)
    it = 0
    for i in range(max_iteration):
        if len(ys):
            ym, wm = ys.mean(0), ws.mean(0)
            H = (ys - ym).T @ (ws - wm)
            U, S, Vt = np.linalg.svd(H)
            D = np.eye(3)
            if np.linalg.det(Vt.T @ U.T) < 0:
                D[2, 2] = -1
            R = Vt.T @ D @ U.T
            upd = np.eye(4)
            upd[:3, :3] = R
            upd[:3, 3] = wm - R @ ym
            T = upd @ T
        nfit, nrm, ys, ws = evaluate(T)
        it = i + 1
        stop = abs(fit - nfit) < relative_fitness and abs(rm - nrm) < relative_rmse
        fit, rm = nfit, nrm
        if stop:
            break
    return T, fit, rm, it
