"""Full-size parity report of the fused solver against fixtures made by the REFERENCE's own
solve_model / predict_model (tests/golden/make_golden.py houv_full*: houv.py:106-206 loaded
unmodified, 2048 points x 64 starts x 200 iterations = BASELINE config 4's shape).

Used by tests/test_full_parity_gpu.py (which asserts on it), by bench.py (`parity_checked`, once,
outside the timed region) and by tools/full_parity.py (writes the table kept under profiles/).
Reads only the committed .npz files -- never oracle/ or /root/reference."""
import os

import numpy as np
import torch

from . import solver
from .loss import rotation_error

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                          "tests", "golden")
CHECKPOINTS = (1, 5, 20, 50, 100, 200)  # forwards (1-based) at which the 64 traces are compared


def _rot_delta_rad(Ra, Rb):
    """angle of Ra Rb^T in radians (float64)"""
    Ra, Rb = np.asarray(Ra, np.float64), np.asarray(Rb, np.float64)
    c = (np.einsum("...ij,...ij->...", Ra, Rb) - 1.0) / 2.0
    return np.arccos(np.clip(c, -1.0, 1.0))


def _call_report(g, c, src, tgt, kernel=64):
    """One predict_model call of the fixture (c = 0: every pair at angle_base 0; c >= 1: the retry
    sub-batch at angle_base c) against one launch of K3 with the same rows."""
    base, npairs = int(g["pm%d_base" % c]), int(g["pm%d_npairs" % c])
    ref_trace = g["pm%d_trace" % c]              # [its, rows]
    its, rows = ref_trace.shape
    params = torch.from_numpy(solver.initial_parameters(rows)).cuda()
    out = solver.pose_optimise(src, tgt, params, kernel, its, angle_base=base, want_trace=True)
    tr = out["trace"][:, :, 1].double().cpu().numpy()
    rel = np.abs(tr - ref_trace) / np.maximum(np.abs(ref_trace), 1e-30)
    rep = dict(angle_base=base, pairs=npairs, rows=rows, iterations=its, checkpoints={})
    for k in CHECKPOINTS:
        if k <= its:
            r = rel[k - 1]
            rep["checkpoints"][str(k)] = dict(
                max_rel=float(r.max()), median_rel=float(np.median(r)),
                frac_within_1e_5=float((r <= 1e-5).mean()), frac_within_1e_4=float((r <= 1e-4).mean()),
                frac_within_1e_3=float((r <= 1e-3).mean()))
    # end state of every start: rotation / translation of the last forward
    R, T = out["R"].cpu().numpy(), out["T"].cpu().numpy()
    Rr, Tr = g["pm%d_R" % c].reshape(rows, 3, 3), g["pm%d_T" % c].reshape(rows, 3)
    dr, dt = _rot_delta_rad(R, Rr), np.linalg.norm(T - Tr, axis=1)
    rep["end_state_all_starts"] = dict(
        rot_rad_median=float(np.median(dr)), rot_rad_max=float(dr.max()),
        trans_median=float(np.median(dt)), trans_max=float(dt.max()),
        frac_within_bar=float(((dr <= 1e-4) & (dt <= 1e-5)).mean()))
    # the start each pair selects (arg-min of min_1 over the 64, houv.py:153)
    m1 = out["min_1"].reshape(npairs, kernel).cpu().numpy()
    m1r = g["pm%d_min1" % c].reshape(npairs, kernel)
    sel, selr = m1.argmin(1), m1r.argmin(1)
    idx = np.arange(npairs)
    rep["selected"] = dict(
        same_start=[bool(a == b) for a, b in zip(sel, selr)],
        best_min_1=[float(x) for x in m1[idx, sel]], best_min_1_ref=[float(x) for x in m1r[idx, selr]],
        rot_rad=[float(x) for x in _rot_delta_rad(R.reshape(npairs, kernel, 3, 3)[idx, sel],
                                                  Rr.reshape(npairs, kernel, 3, 3)[idx, selr])],
        trans=[float(x) for x in np.linalg.norm(T.reshape(npairs, kernel, 3)[idx, sel] -
                                                Tr.reshape(npairs, kernel, 3)[idx, selr], axis=1)])
    return rep, m1r


def reference_envelope(name="houv_full", alt="houv_full_alt", kernel=64):
    """The reference against ITSELF: `alt` is the same run of the reference's solve_model with its
    gradient kernel adding in the opposite point order (the reference uses float atomics, whose
    order the GPU is free to choose: chamfer3D.cu:166-171, registration/README.md:82-90).  Same
    statistics as _call_report -- the yardstick for what 'parity' can mean at 200 iterations."""
    a = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    b = np.load(os.path.join(GOLDEN_DIR, alt + ".npz"))
    ta, tb = a["pm0_trace"], b["pm0_trace"]
    rel = np.abs(tb - ta) / np.maximum(np.abs(ta), 1e-30)
    rep = dict(fixtures=[name + ".npz", alt + ".npz"], checkpoints={})
    for k in CHECKPOINTS:
        r = rel[k - 1]
        rep["checkpoints"][str(k)] = dict(
            max_rel=float(r.max()), median_rel=float(np.median(r)),
            frac_within_1e_5=float((r <= 1e-5).mean()), frac_within_1e_4=float((r <= 1e-4).mean()),
            frac_within_1e_3=float((r <= 1e-3).mean()))
    rows = ta.shape[1]
    npairs = rows // kernel
    Ra, Rb = a["pm0_R"].reshape(rows, 3, 3), b["pm0_R"].reshape(rows, 3, 3)
    Ta, Tb = a["pm0_T"].reshape(rows, 3), b["pm0_T"].reshape(rows, 3)
    dr, dt = _rot_delta_rad(Ra, Rb), np.linalg.norm(Ta - Tb, axis=1)
    rep["end_state_all_starts"] = dict(
        rot_rad_median=float(np.median(dr)), rot_rad_max=float(dr.max()),
        trans_median=float(np.median(dt)), trans_max=float(dt.max()),
        frac_within_bar=float(((dr <= 1e-4) & (dt <= 1e-5)).mean()))
    sa, sb = a["pm0_min1"].argmin(1), b["pm0_min1"].argmin(1)
    idx = np.arange(npairs)
    rep["selected"] = dict(
        same_start=[bool(x == y) for x, y in zip(sa, sb)],
        rot_rad=[float(x) for x in _rot_delta_rad(a["pm0_R"][idx, sa], b["pm0_R"][idx, sb])],
        trans=[float(x) for x in np.linalg.norm(a["pm0_T"][idx, sa] - b["pm0_T"][idx, sb], axis=1)],
        best_min_1_rel=[float(abs(x - y) / x) for x, y in zip(a["pm0_min1"][idx, sa],
                                                              b["pm0_min1"][idx, sb])])
    rep["metrics"] = dict(r_err_deg=[[float(x) for x in a["r_err"]], [float(x) for x in b["r_err"]]],
                          t_err=[[float(x) for x in a["t_err"]], [float(x) for x in b["t_err"]]])
    return rep


def full_parity_report(name="houv_full", calls=None):
    """-> dict: per predict_model call the trace deviations at CHECKPOINTS, end states, selected
    starts; and solve_model's ans / metrics against the reference's."""
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    src, tgt = torch.from_numpy(g["src"]).cuda(), torch.from_numpy(g["tgt"]).cuda()
    gt = torch.from_numpy(g["transform"]).cuda()
    n_calls = int(g["n_calls"])
    rep = dict(fixture=name + ".npz", shape="%d pairs x 64 starts x %d its x %d pts" % (
        src.shape[0], g["pm0_trace"].shape[0], src.shape[1]), calls=[])
    r0, m1r = _call_report(g, 0, src, tgt)
    rep["calls"].append(r0)
    if n_calls > 1 and (calls is None or calls > 1):
        # the reference's retry sub-batch: pairs whose best start stays above 0.030 (houv.py:154-157)
        lst = np.nonzero(m1r.min(1) > solver.RETRY_THRESHOLD)[0]
        assert len(lst) == int(g["pm1_npairs"])
        for c in range(1, n_calls):
            rep["calls"].append(_call_report(g, c, src[lst].contiguous(), tgt[lst].contiguous())[0])
        rep["retried_pairs_ref"] = [int(i) for i in lst]
    # the public call
    r_err, t_err, ans = solver.solve_model(solver.HOUV(64, 0), src, tgt, gt, kernel=64,
                                           num_epochs=g["pm0_trace"].shape[0])
    a, ar = ans.cpu().numpy(), g["ans"]
    rep["solve_model"] = dict(
        rot_rad=[float(x) for x in _rot_delta_rad(a[:, :3, :3], ar[:, :3, :3])],
        trans=[float(x) for x in np.linalg.norm(a[:, :3, 3] - ar[:, :3, 3], axis=1)],
        r_err_deg=[float(x) for x in r_err.cpu()], r_err_deg_ref=[float(x) for x in g["r_err"]],
        t_err=[float(x) for x in t_err.cpu()], t_err_ref=[float(x) for x in g["t_err"]],
        bottom_row_zero=bool((a[:, 3] == 0).all()))
    if os.path.exists(os.path.join(GOLDEN_DIR, name + "_alt.npz")):
        rep["reference_against_itself"] = reference_envelope(name, name + "_alt")
    return rep
