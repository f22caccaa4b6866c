"""Drop-in for the reference's HOUV solver module
(/root/reference/registration/models/houv.py): same names, signatures, return
values and error behaviour for
    HOUV (:13-103), predict_model (:106-138), solve_model (:142-206), Predict_loss (:209-222)
so that train_HOUV.py:38,84,112 runs by changing only the import.

predict_model's 200-iteration loop is ONE launch of the fused K3 kernel
(houv_pose_optimise) instead of ~120 launches per iteration; HOUV.forward and
Predict_loss remain available as plain torch code over the K1/K2 operator for
callers that build their own loop (and as the in-repo cross-check of K3).
"""
import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .loss import Predict_loss, rotation_error, translation_error  # noqa: F401 (re-export)

ADAM_LR = 0.01          # houv.py:118
ADAM_BETAS = (0.9, 0.999)
ADAM_EPS = 1e-8
RETRY_THRESHOLD = 0.030  # houv.py:156
W_3D, W_VIEW, ALPHA = 6.0, 1.0, 0.5  # houv.py:209,222

# When set to a list, every K3 launch appends (start_event, end_event, rows, iters, N)
# recorded on the launching stream: bench.py derives the kernel's roofline from it.
KERNEL_TIMING = None
FUSED_MAX_POINTS = 2048  # csrc/pose_common.cuh kPts: both clouds of a pair live in one SM's shared memory
# What the last solve_model_batches call did on this rank: pairs, pairs sent through the retry
# cascade (best min_1 above 0.030 after angle_base 0), rows of its first-pass and retry launches
# (the retry rows this rank RAN, which differ from the ones it owns when rows are dealt over ranks).
LAST_STATS = {}


def initial_parameters(rows, seed=2021):
    """The draws of HOUV.reset_weight (houv.py:40-61) as one [rows, 8] float32
    array (V_c[3], angle_c, tran_c[3], tran_s).  NumPy is re-seeded before every
    tensor, so angle_c == tran_s and tran_c reuses V_c's stream; rows 0..25 of
    V_c are overwritten with the axis set {-1,0,1}^3 minus 0.  Row i's values do
    not depend on `rows` (row-major fill of one seeded stream)."""
    np.random.seed(seed)
    vc = np.random.randn(rows, 3)
    num = 0
    for x0 in [-1, 0, 1]:
        for y0 in [-1, 0, 1]:
            for z0 in [-1, 0, 1]:
                if x0 == 0 and y0 == 0 and z0 == 0:
                    continue
                vc[num] = np.array((x0, y0, z0))  # IndexError when rows < 26, as the reference
                num += 1
    np.random.seed(seed)
    angle_c = np.random.randn(rows, 1)
    np.random.seed(seed)
    tran_c = np.random.randn(rows, 3)
    np.random.seed(seed)
    tran_s = np.random.randn(rows, 1)
    return np.concatenate([vc, angle_c, tran_c, tran_s], axis=1).astype(np.float32)


class HOUV(nn.Module):
    """houv.py:13-103.  Parameters live on the GPU; `forward` is the reference's
    torch expression (autograd-capable)."""

    def __init__(self, batch_size, angle_base, device="cuda"):
        super(HOUV, self).__init__()
        self.batch_size = batch_size
        self.angle_base = angle_base
        self.pi = torch.acos(torch.zeros(1)).item() * 2
        self._device = torch.device(device)
        # the reference draws unseeded values here (houv.py:21-36) that reset_weight always
        # replaces before use; we keep the shapes only
        z = lambda c: nn.Parameter(torch.zeros(batch_size, c, device=self._device))
        self.V_c, self.angle_c, self.tran_c, self.tran_s = z(3), z(1), z(3), z(1)

    def set_packed(self, packed):
        """Adopt a [rows,8] device tensor of parameters (V_c, angle_c, tran_c, tran_s)."""
        self.batch_size = packed.shape[0]
        self.V_c = nn.Parameter(packed[:, 0:3].contiguous())
        self.angle_c = nn.Parameter(packed[:, 3:4].contiguous())
        self.tran_c = nn.Parameter(packed[:, 4:7].contiguous())
        self.tran_s = nn.Parameter(packed[:, 7:8].contiguous())

    def packed(self):
        return torch.cat([self.V_c, self.angle_c, self.tran_c, self.tran_s], dim=1).detach()

    def reset_weight(self, batch_size, angle_base, seed=2021):
        self.angle_base = angle_base
        host = torch.from_numpy(initial_parameters(batch_size, seed))
        self.set_packed(host.to(self._device))
        self.ones = torch.zeros((batch_size, 3, 3), device=self._device) + \
            torch.eye(3, device=self._device)

    def cd_rotation(self, angle, V, device='cuda'):
        batch_size = angle.shape[0]
        V = V / torch.sqrt((V * V).sum(dim=1, keepdim=True))
        A = torch.zeros((batch_size, 3, 3), dtype=torch.float, device=V.device)
        A[:, 0, 1] = -V[:, 2]
        A[:, 0, 2] = V[:, 1]
        A[:, 1, 0] = V[:, 2]
        A[:, 1, 2] = -V[:, 0]
        A[:, 2, 0] = -V[:, 1]
        A[:, 2, 1] = V[:, 0]
        ones = torch.eye(3, device=V.device).expand(batch_size, 3, 3)
        R = ones + torch.sin(angle).unsqueeze(2) * A + \
            (1 - torch.cos(angle)).unsqueeze(2) * torch.bmm(A, A)
        return R

    def translation(self, tran, s):
        tran = tran / torch.sqrt((tran * tran).sum(dim=1, keepdim=True))
        tran = tran * s
        tran = tran.unsqueeze(1)
        return tran

    def forward(self, src):
        src = src.squeeze(0)
        angle = torch.sin(self.angle_c * self.pi) * self.pi / 8 + self.pi / 8 + \
            self.angle_base * self.pi / 4
        R = self.cd_rotation(angle, self.V_c)
        tran_s = torch.sin(self.tran_s * self.pi) * 0.125 + 0.125
        T = self.translation(self.tran_c, tran_s)
        src_temp = torch.bmm(src, R.transpose(1, 2)) + T
        return src_temp, R, T


def _pose_optimise_large(src, tgt, params, kernel, iters, angle_base, grad_scale, adam_m, adam_v,
                         step0, lr, betas, eps, percent, w3d, wview, want_loss, want_trace,
                         row_pair, row_base, tran_scale, tran_offset, want_tran_s, row_scale,
                         rows_per_chunk=64):
    """pose_optimise for clouds of more than FUSED_MAX_POINTS points, which the fused kernel cannot
    keep resident: the reference's own scheme (houv.py:94-126 -- torch forward, Predict_loss terms,
    autograd, Adam) on the device over THIS package's Chamfer operator (K1 / K2 take any size,
    chamfer3D.cu:15,22 grid-stride semantics), rows in chunks.  Same contract as the fused path:
    per-row pair / angle_base / gradient scale, float32 or float64 state updated in place, resume
    through step0 / adam_m / adam_v, results of the LAST forward.  The k-th-rank tie rule is
    torch.topk's here."""
    from . import loss as L
    dev, rows = src.device, params.shape[0]
    P, N, _ = src.shape
    if row_pair is None:
        row_pair = torch.arange(P, device=dev, dtype=torch.int32).repeat_interleave(kernel)
    pair = row_pair.long()
    base = (row_base.float() if row_base is not None
            else torch.full((rows,), float(angle_base), device=dev))
    scale = (row_scale if row_scale is not None
             else torch.full((rows,), float(grad_scale), device=dev))
    pi = torch.acos(torch.zeros(1)).item() * 2                      # float32 pi, houv.py:19
    m = adam_m if adam_m is not None else torch.zeros_like(params)
    v = adam_v if adam_v is not None else torch.zeros_like(params)
    min1 = torch.empty(rows, device=dev)
    R_out, T_out = torch.empty(rows, 3, 3, device=dev), torch.empty(rows, 3, device=dev)
    loss_out = torch.empty(rows, device=dev) if want_loss else None
    ts_out = torch.empty(rows, device=dev) if want_tran_s else None
    trace = torch.empty(iters, rows, 2, device=dev) if want_trace else None
    eye = torch.eye(3, device=dev)
    for it in range(iters):
        grad = torch.zeros(rows, 8, device=dev)
        for lo in range(0, rows, rows_per_chunk):
            sl = slice(lo, min(lo + rows_per_chunk, rows))
            p = params[sl].float().detach().requires_grad_(True)    # float64 leaves feed float32 (train_utils.py:397-401)
            s_rows, t_rows = src[pair[sl]], tgt[pair[sl]]
            V = p[:, 0:3] / torch.sqrt((p[:, 0:3] * p[:, 0:3]).sum(1, keepdim=True))
            ang = torch.sin(p[:, 3:4] * pi) * pi / 8 + pi / 8 + base[sl].unsqueeze(1) * pi / 4
            A = torch.zeros(p.shape[0], 3, 3, device=dev)
            A[:, 0, 1], A[:, 0, 2], A[:, 1, 0] = -V[:, 2], V[:, 1], V[:, 2]
            A[:, 1, 2], A[:, 2, 0], A[:, 2, 1] = -V[:, 0], -V[:, 1], V[:, 0]
            Rm = eye + torch.sin(ang).unsqueeze(2) * A + (1 - torch.cos(ang)).unsqueeze(2) * torch.bmm(A, A)
            tsz = torch.sin(p[:, 7:8] * pi) * tran_scale + tran_offset
            Tv = (p[:, 4:7] / torch.sqrt((p[:, 4:7] * p[:, 4:7]).sum(1, keepdim=True)) * tsz).unsqueeze(1)
            y = torch.bmm(s_rows, Rm.transpose(1, 2)) + Tv
            cd_t, cd_p = L.calc_cd_percent(y, t_rows, percent=percent)
            min_1 = torch.min(torch.stack([cd_t, cd_p], 1), dim=1).values
            loss = min_1 * w3d
            if wview != 0.0:
                for dim in range(3):
                    a, b = L.loss_view(y, t_rows, dim=dim)
                    loss = loss + wview * torch.min(torch.stack([a, b], 1), dim=1).values
            (loss * scale[sl]).sum().backward()
            grad[sl] = p.grad
            if trace is not None:
                trace[it, sl, 0], trace[it, sl, 1] = loss.detach(), min_1.detach()
            if it == iters - 1:
                min1[sl], R_out[sl], T_out[sl] = min_1.detach(), Rm.detach(), Tv.detach()[:, 0]
                if loss_out is not None:
                    loss_out[sl] = loss.detach()
                if ts_out is not None:
                    ts_out[sl] = tsz.detach()[:, 0]
        # torch.optim.Adam (no weight decay / amsgrad), in the precision of the state
        t = step0 + it + 1
        g = grad.to(params.dtype)
        m.mul_(betas[0]).add_(g, alpha=1 - betas[0])
        v.mul_(betas[1]).addcmul_(g, g, value=1 - betas[1])
        denom = v.sqrt() / (1 - betas[1] ** t) ** 0.5 + eps
        params.addcdiv_(m, denom, value=-lr / (1 - betas[0] ** t))
    return dict(min_1=min1, R=R_out, T=T_out, loss=loss_out, trace=trace, tran_s=ts_out)


def _k3_kernel_launches(rows, kernel, has_row_pair, wview, dev):
    """Kernels one houv_pose_optimise_ex call launches (for bench.py's gpu_launches): mirrors
    csrc/pose_duo.cu launch_duo for the default variant -- whole waves of two-row CTAs, the rows that
    cannot be paired (caller-made row_pair, odd K) through the one-row kernel, and a one-row tail when
    what is left fits one wave.  Other variants (HOUV_POSE_VARIANT, a loss without view terms) launch one."""
    import os
    if wview == 0.0 or os.environ.get("HOUV_POSE_VARIANT", "20") not in ("20", "21") or rows == 0:
        return 1 if rows else 0
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    nduo = rows // 2
    whole = nduo // sms * sms
    if whole > 0 and rows - 2 * whole <= sms:
        nduo = whole
    n = 0
    if nduo > 0:
        n += 1 + (1 if (has_row_pair or kernel % 2) else 0)
    if rows > 2 * nduo:
        n += 1
    return n


def pose_optimise(src, tgt, params, kernel, iters, angle_base=0, grad_scale=None,
                  adam_m=None, adam_v=None, step0=0, lr=ADAM_LR, betas=ADAM_BETAS, eps=ADAM_EPS,
                  percent=ALPHA, w3d=W_3D, wview=W_VIEW, want_loss=False, want_trace=False,
                  row_pair=None, row_base=None, tran_scale=0.125, tran_offset=0.125,
                  want_tran_s=False, row_scale=None):
    """Thin wrapper over the C ABI houv_pose_optimise_ex (K3).
    src[P,N,3], tgt[P,N,3] device float32; params[P*kernel,8] is updated IN PLACE: float32
    (HOUV, houv.py) or float64 (train_utils.getPredict_angle keeps float64 leaves; adam_m /
    adam_v must then be float64 too).
    row_pair / row_base (int32 device tensors [rows]) optionally map every row to a pair
    and an angle_base of its own; rows is then params.shape[0] instead of P*kernel.
    row_scale (float32 [rows]) gives every row its own grad_scale (several batches in one launch).
    -> dict(min_1[rows], R[rows,3,3], T[rows,3], loss[rows]|None, trace[iters,rows,2]|None,
            tran_s[rows]|None)"""
    _lib.require_cuda_f32(src, "src", 3)
    _lib.require_cuda_f32(tgt, "tgt", 3)
    if not isinstance(params, torch.Tensor) or not params.is_cuda or \
            params.dtype not in (torch.float32, torch.float64):
        raise RuntimeError("params must be a float32 or float64 CUDA tensor "
                           "(houv_b200 has no CPU path)")
    if iters <= 0:
        # the reference's loop would leave min_1 / R / T undefined (houv.py:134 NameError)
        raise RuntimeError("pose_optimise needs iters >= 1 (got %d)" % iters)
    f64 = params.dtype == torch.float64
    src, tgt = src.contiguous(), tgt.contiguous()
    P, N, _ = src.shape
    M = tgt.shape[1]
    rows = P * kernel if row_pair is None else int(row_pair.shape[0])
    if tuple(params.shape) != (rows, 8) or not params.is_contiguous():
        raise RuntimeError("params must be a contiguous [rows, 8] tensor (rows = P*kernel)")
    for name, a in (("adam_m", adam_m), ("adam_v", adam_v)):
        if a is not None and (a.dtype != params.dtype or not a.is_cuda or
                              tuple(a.shape) != (rows, 8) or not a.is_contiguous()):
            raise RuntimeError(name + " must match params in dtype, device and shape")
    for name, a in (("row_pair", row_pair), ("row_base", row_base)):
        if a is not None and (a.dtype != torch.int32 or not a.is_cuda or a.numel() != rows
                              or not a.is_contiguous()):
            raise RuntimeError(name + " must be a contiguous int32 CUDA tensor of [rows]")
    if tgt.shape[0] != P:
        raise RuntimeError("src and tgt hold different numbers of pairs")
    dev = src.device
    min1 = torch.empty(rows, device=dev)
    R = torch.empty(rows, 3, 3, device=dev)
    T = torch.empty(rows, 3, device=dev)
    loss = torch.empty(rows, device=dev) if want_loss else None
    trace = torch.empty(iters, rows, 2, device=dev) if want_trace else None
    tran_s = torch.empty(rows, device=dev) if want_tran_s else None
    ptr = lambda x: 0 if x is None else x.data_ptr()
    if grad_scale is None:
        grad_scale = 1.0 / rows  # rmse_loss.mean() over every row of the batch, houv.py:124
    if row_scale is not None and (row_scale.dtype != torch.float32 or not row_scale.is_cuda
                                  or row_scale.numel() != rows or not row_scale.is_contiguous()):
        raise RuntimeError("row_scale must be a contiguous float32 CUDA tensor of [rows]")
    if max(N, M) > FUSED_MAX_POINTS:
        # the reference's loop takes any cloud size (chamfer3D.cu:15,22); so does this call
        if N != M:
            raise RuntimeError("src and tgt must hold the same number of points "
                               "(calc_cd_percent trims both lists to k = N*percent)")
        return _pose_optimise_large(src, tgt, params, kernel, iters, angle_base, grad_scale, adam_m,
                                    adam_v, step0, lr, betas, eps, percent, w3d, wview, want_loss,
                                    want_trace, row_pair, row_base, tran_scale, tran_offset,
                                    want_tran_s, row_scale)
    q = _lib.HouvPoseProblem()
    q.src, q.tgt, q.P, q.N, q.M, q.K = src.data_ptr(), tgt.data_ptr(), P, N, M, kernel
    if f64:
        q.params64, q.adam_m64, q.adam_v64 = params.data_ptr(), ptr(adam_m), ptr(adam_v)
    else:
        q.params, q.adam_m, q.adam_v = params.data_ptr(), ptr(adam_m), ptr(adam_v)
    q.step0, q.iters, q.angle_base = step0, iters, int(angle_base)
    q.lr, q.beta1, q.beta2, q.eps = lr, betas[0], betas[1], eps
    q.grad_scale, q.percent, q.w3d, q.wview = grad_scale, percent, w3d, wview
    q.tran_scale, q.tran_offset = tran_scale, tran_offset
    q.min1_out, q.loss_out, q.R_out, q.T_out = min1.data_ptr(), ptr(loss), R.data_ptr(), T.data_ptr()
    q.tran_s_out, q.trace = ptr(tran_s), ptr(trace)
    q.row_pair, q.row_base, q.rows = ptr(row_pair), ptr(row_base), rows
    q.row_scale = ptr(row_scale)
    timing = KERNEL_TIMING
    if timing is not None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
    import ctypes
    with torch.cuda.device(dev):
        rc = _lib.load().houv_pose_optimise_ex(ctypes.byref(q), _lib.stream_ptr())
    _lib.check(rc, "houv_pose_optimise_ex")
    _lib.count_launches(_k3_kernel_launches(rows, kernel, row_pair is not None, wview, dev))
    if timing is not None:
        ev1.record()
        timing.append((ev0, ev1, rows, iters, N))
    return dict(min_1=min1, R=R, T=T, loss=loss, trace=trace, tran_s=tran_s)


def predict_model(net, src, src_rotated, pose=None, src_ori=None, tgt_ori=None, angle_t=None,
                  label=None, kernel=64, num_epochs=500, angle_base=0, device='cuda', seed=2021):
    """houv.py:106-138: every pair is optimised from `kernel` starts for num_epochs
    Adam iterations; returns (min_1[B,kernel], R[B,kernel,3,3], T[B,kernel,3]) of the
    LAST forward.  `net` ends up holding the parameters after num_epochs steps."""
    batch_size = src.shape[0]
    rows = batch_size * kernel
    net.reset_weight(rows, angle_base, seed=seed)
    params = net.packed().contiguous()
    out = pose_optimise(src.float(), src_rotated.float(), params, kernel, num_epochs,
                        angle_base=angle_base)
    net.set_packed(params)
    rmse_loss = out["min_1"].reshape((batch_size, kernel))
    R = out["R"].reshape((batch_size, kernel, 3, 3))
    T = out["T"].reshape((batch_size, kernel, 3))
    return rmse_loss, R, T


def solve_model(net, src, src_rotated, pose=None, src_ori=None, tgt_ori=None, angle_t=None,
                label=None, kernel=64, num_epochs=200, prefix='train'):
    """houv.py:142-206: angle_base 0 for every pair, then angle_base 1..3 for the pairs
    whose best start stays above 0.030, keeping a retry when it improves the pair's best;
    ans[B,4,4] carries R and T of the best start and an all-zero bottom row (:187-195)."""
    batch_size = src.shape[0]
    rmse_loss, R, T = predict_model(net, src, src_rotated, pose, src_ori, tgt_ori, angle_t, label,
                                    kernel=kernel, num_epochs=num_epochs, angle_base=0)
    rmse, b_index = rmse_loss.topk(1, dim=1, largest=False, sorted=True)
    lst_add = torch.nonzero(rmse[:, 0] > RETRY_THRESHOLD).reshape(-1)
    n_add = int(lst_add.numel())
    if n_add > 0:
        # The reference runs predict_model three more times on the sub-batch (angle_base
        # 1, 2, 3; houv.py:168-169).  The three runs are independent of one another (same
        # seeded initial values, same 1/(n_add*kernel) gradient scale), so they go out as ONE
        # launch of 3*n_add*kernel rows; the keep-if-better updates are then applied in the
        # reference's order (:171-180).
        sub_rows = n_add * kernel
        net.reset_weight(sub_rows, 3, seed=2021)
        params = net.packed().repeat(3, 1).contiguous()
        row_pair = torch.arange(n_add, device=src.device, dtype=torch.int32) \
            .repeat_interleave(kernel).repeat(3).contiguous()
        row_base = torch.arange(1, 4, device=src.device, dtype=torch.int32) \
            .repeat_interleave(sub_rows).contiguous()
        out = pose_optimise(src[lst_add].float(), src_rotated[lst_add].float(), params, kernel,
                            num_epochs, grad_scale=1.0 / sub_rows, row_pair=row_pair,
                            row_base=row_base)
        net.set_packed(params[2 * sub_rows:])
        for b in range(3):
            sl = slice(b * sub_rows, (b + 1) * sub_rows)
            rmse_loss_add = out["min_1"][sl].reshape(n_add, kernel)
            R_add = out["R"][sl].reshape(n_add, kernel, 3, 3)
            T_add = out["T"][sl].reshape(n_add, kernel, 3)
            rmse_add, _ = rmse_loss_add.topk(1, dim=1, largest=False, sorted=True)
            flag = torch.nonzero((rmse_add < rmse[lst_add]).reshape(-1)).reshape(-1)
            geid = lst_add[flag]
            R[geid] = R_add[flag]
            rmse_loss[geid] = rmse_loss_add[flag]
            T[geid] = T_add[flag]
            rmse[geid] = rmse_add[flag]
    best = rmse_loss.argmin(dim=1)  # first minimum, like topk(1, largest=False)
    rows = torch.arange(batch_size, device=src.device)
    ans = torch.zeros((batch_size, 4, 4), device=src.device)
    ans[:, :3, :3] = R[rows, best]
    ans[:, :3, 3] = T[rows, best]
    if prefix == 'test':
        return ans.cpu()
    if prefix == 'device':  # extension: same as 'test' without the device->host copy
        return ans
    r_err = rotation_error(ans[:, :3, :3], pose[:, :3, :3])
    t_err = translation_error(ans[:, :3, 3], pose[:, :3, 3])
    print("Rotation error:", r_err.mean(), "Translation error:", t_err.mean())
    return r_err, t_err, ans


def _balanced_retry(srcF, tgtF, p_rows, pair_rows, base_rows, scale_rows, kernel, num_epochs):
    """The retry launch of solve_model_batches with the rows of every rank dealt evenly over the
    ranks (pairs are independent and so are rows: a row is its pair's two clouds, its initial
    values, angle_base and gradient scale).  Three small all-gathers (NCCL): failed-pair counts +
    clouds, row descriptions, row results.  -> dict(min_1, R, T) of THIS rank's rows, the same
    bits pose_optimise would give for them locally."""
    dist = torch.distributed
    world, rank, dev = dist.get_world_size(), dist.get_rank(), srcF.device
    n_l, N = srcF.shape[0], srcF.shape[1]
    rows_l = sum(int(p.shape[0]) for p in p_rows)
    cnt = torch.tensor([n_l, rows_l], device=dev, dtype=torch.int64)
    cnts = torch.empty(world * 2, device=dev, dtype=torch.int64)  # outputs are dim-0 concatenations
    dist.all_gather_into_tensor(cnts, cnt)                        # (the form gloo accepts as well as NCCL)
    cnts = cnts.reshape(world, 2).cpu()
    n_all, r_all = cnts[:, 0].tolist(), cnts[:, 1].tolist()
    nmax, rmax, r_tot = max(n_all), max(r_all), sum(r_all)
    empty = dict(min_1=torch.zeros(0, device=dev), R=torch.zeros(0, 3, 3, device=dev),
                 T=torch.zeros(0, 3, device=dev))
    if r_tot == 0:
        return empty
    # failed pairs of every rank, padded to nmax per rank: global pair index = rank*nmax + local
    clouds = torch.zeros(2, nmax, N, 3, device=dev)
    clouds[0, :n_l], clouds[1, :n_l] = srcF, tgtF
    g_clouds = torch.empty(world * 2, nmax, N, 3, device=dev)
    dist.all_gather_into_tensor(g_clouds, clouds)
    g_clouds = g_clouds.reshape(world, 2, nmax, N, 3)
    g_src = g_clouds[:, 0].reshape(world * nmax, N, 3).contiguous()
    g_tgt = g_clouds[:, 1].reshape(world * nmax, N, 3).contiguous()
    # row descriptions [rmax, 11]: 8 initial values, pair (global), angle_base, gradient scale
    desc = torch.zeros(rmax, 11, device=dev)
    if rows_l:
        desc[:rows_l, :8] = torch.cat(p_rows)
        desc[:rows_l, 8] = (torch.cat(pair_rows) + rank * nmax).float()  # < 2^24: exact in float32
        desc[:rows_l, 9] = torch.cat(base_rows).float()
        desc[:rows_l, 10] = torch.cat(scale_rows)
    g_desc = torch.empty(world * rmax, 11, device=dev)
    dist.all_gather_into_tensor(g_desc, desc)
    g_desc = g_desc.reshape(world, rmax, 11)
    rows = torch.cat([g_desc[k, :r_all[k]] for k in range(world)])  # [r_tot, 11], rank-major
    chunk = (r_tot + world - 1) // world
    lo, hi = min(rank * chunk, r_tot), min((rank + 1) * chunk, r_tot)
    res = torch.zeros(chunk, 13, device=dev)
    if hi > lo:
        mine = rows[lo:hi]
        o = pose_optimise(g_src, g_tgt, mine[:, :8].contiguous(), kernel, num_epochs,
                          row_pair=mine[:, 8].to(torch.int32).contiguous(),
                          row_base=mine[:, 9].to(torch.int32).contiguous(),
                          row_scale=mine[:, 10].contiguous())
        res[:hi - lo, 0] = o["min_1"]
        res[:hi - lo, 1:10] = o["R"].reshape(-1, 9)
        res[:hi - lo, 10:13] = o["T"]
    g_res = torch.empty(world * chunk, 13, device=dev)
    dist.all_gather_into_tensor(g_res, res)  # rank k's rows [k*chunk, ...) = global rows in order
    if rows_l == 0:
        return empty
    off = sum(r_all[:rank])
    own = g_res[off:off + rows_l]
    return dict(min_1=own[:, 0].contiguous(), R=own[:, 1:10].reshape(-1, 3, 3).contiguous(),
                T=own[:, 10:13].contiguous())


def solve_model_batches(src, src_rotated, batch_size=50, kernel=64, num_epochs=200, balance=None):
    """solve_model (houv.py:142-206) for a whole set of pairs walked in batches of `batch_size`
    as train_HOUV.py:100-112 does, with the SAME per-pair results as calling solve_model batch by
    batch -- initial values by position in the batch (houv.py:116), gradient scale 1/(rows of
    the batch) (:124), retries at angle_base 1..3 on each batch's own sub-batch (:154-180) -- but
    in TWO launches for the whole set (all first passes, then all retries) instead of two per
    batch, which removes the partial last wave of every batch.  -> ans[P,4,4] on the device.
    `balance` (default: on when torch.distributed runs with more than one rank; every rank must
    then call this together): the retry rows of ALL ranks are dealt evenly over the ranks before
    the retry launch (_balanced_retry) -- same rows, same results, no rank waits for the one
    whose shard has the most failed pairs."""
    src, src_rotated = src.float().contiguous(), src_rotated.float().contiguous()
    P, dev = src.shape[0], src.device
    if balance is None:
        balance = torch.distributed.is_available() and torch.distributed.is_initialized() \
            and torch.distributed.get_world_size() > 1
    if P == 0:
        if balance:  # a rank without pairs still takes its share of the other ranks' retry rows
            _balanced_retry(src, src_rotated, [], [], [], [], kernel, num_epochs)
        return torch.zeros(0, 4, 4, device=dev)
    bounds = [(l, min(l + batch_size, P)) for l in range(0, P, batch_size)]
    # row i of reset_weight's draws does not depend on the number of rows (initial_parameters)
    init = torch.from_numpy(initial_parameters(min(batch_size, P) * kernel)).to(dev)
    params = torch.cat([init[:(r - l) * kernel] for l, r in bounds]).contiguous()
    scale = torch.cat([torch.full(((r - l) * kernel,), 1.0 / ((r - l) * kernel), device=dev)
                       for l, r in bounds]).contiguous()
    out = pose_optimise(src, src_rotated, params, kernel, num_epochs, angle_base=0, row_scale=scale)
    rmse_loss = out["min_1"].reshape(P, kernel)
    R = out["R"].reshape(P, kernel, 3, 3)
    T = out["T"].reshape(P, kernel, 3)
    rmse = rmse_loss.min(dim=1).values
    lst_add = torch.nonzero(rmse > RETRY_THRESHOLD).reshape(-1)  # ascending = batch order
    n_add = int(lst_add.numel())
    LAST_STATS.clear()
    LAST_STATS.update(pairs=P, retried_pairs=n_add, rows_first=P * kernel,
                      rows_retry_owned=3 * n_add * kernel)
    if n_add > 0 or balance:
        batch_of = (lst_add // batch_size).cpu()
        counts = torch.bincount(batch_of, minlength=len(bounds)).tolist()
        p_rows, pair_rows, base_rows, scale_rows, first = [], [], [], [], 0
        for nb in counts:  # rows laid out [batch][angle_base][pair][start]
            if nb == 0:
                continue
            sub_rows = nb * kernel
            local = torch.arange(first, first + nb, device=dev, dtype=torch.int32) \
                .repeat_interleave(kernel)
            for b in (1, 2, 3):
                p_rows.append(init[:sub_rows])
                pair_rows.append(local)
                base_rows.append(torch.full((sub_rows,), b, device=dev, dtype=torch.int32))
                scale_rows.append(torch.full((sub_rows,), 1.0 / sub_rows, device=dev))
            first += nb
        if balance:
            out2 = _balanced_retry(src[lst_add], src_rotated[lst_add], p_rows, pair_rows, base_rows,
                                   scale_rows, kernel, num_epochs)
        else:
            out2 = pose_optimise(src[lst_add], src_rotated[lst_add], torch.cat(p_rows).contiguous(),
                                 kernel, num_epochs, row_pair=torch.cat(pair_rows).contiguous(),
                                 row_base=torch.cat(base_rows).contiguous(),
                                 row_scale=torch.cat(scale_rows).contiguous())
        # back to [angle_base][pair] order, then the reference's keep-if-better sequence
        order, off, first = [[], [], []], 0, 0
        for nb in counts:
            for b in range(3):
                order[b].append(torch.arange(off, off + nb * kernel, device=dev))
                off += nb * kernel
        for b in range(3 if n_add > 0 else 0):
            sel = torch.cat(order[b])
            l_add = out2["min_1"][sel].reshape(n_add, kernel)
            R_add = out2["R"][sel].reshape(n_add, kernel, 3, 3)
            T_add = out2["T"][sel].reshape(n_add, kernel, 3)
            rmse_add = l_add.min(dim=1).values
            flag = torch.nonzero(rmse_add < rmse[lst_add]).reshape(-1)
            geid = lst_add[flag]
            R[geid] = R_add[flag]
            rmse_loss[geid] = l_add[flag]
            T[geid] = T_add[flag]
            rmse[geid] = rmse_add[flag]
    best = rmse_loss.argmin(dim=1)
    rows = torch.arange(P, device=dev)
    ans = torch.zeros((P, 4, 4), device=dev)
    ans[:, :3, :3] = R[rows, best]
    ans[:, :3, 3] = T[rows, best]
    return ans
