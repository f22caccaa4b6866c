"""Distributional parity (SURVEY §7 hard part 1, report (iv)): the fused kernel against the
reference's own GPU loop -- torch forward / autograd / Adam over the reference's unmodified CUDA
extension (tests/refloop.py) -- on the same B200, 16 synthetic MVP-shaped pairs x 64 starts x 200
iterations at 2048 points.  The reference's float atomics make ITS result order-dependent
(registration/README.md:82-90), so the comparison is per pair on the selected start and over the
set on the metrics, not per trajectory.  Prints the table kept in profiles/r2_distribution_parity.json."""
import json

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def distribution_report(first=0, pairs=16, iters=200):
    from houv_b200 import loss as L, solver, synth
    from tests.refloop import reference_predict_model
    s, t, tr = synth.make_pairs(first, pairs)
    src, tgt, gt = (torch.from_numpy(x).cuda() for x in (s, t, tr))
    m_ref, R_ref, T_ref = reference_predict_model(src, tgt, num_epochs=iters)
    # a second execution of the reference loop: its float atomics are free to land in another order
    m_ref2, _, _ = reference_predict_model(src, tgt, num_epochs=iters)
    m_our, R_our, T_our = solver.predict_model(solver.HOUV(pairs * 64, 0), src, tgt, kernel=64,
                                               num_epochs=iters)
    idx = torch.arange(pairs, device=src.device)
    rep = dict(pairs=pairs, iterations=iters, points=int(src.shape[1]))
    side = {}
    for name, (m, R, T) in dict(reference_loop=(m_ref, R_ref, T_ref), fused=(m_our, R_our, T_our)).items():
        sel = m.argmin(1)
        ans = torch.zeros(pairs, 4, 4, device=src.device)
        ans[:, :3, :3], ans[:, :3, 3] = R[idx, sel], T[idx, sel]
        re = L.rotation_error(ans[:, :3, :3], gt[:, :3, :3])
        te = L.translation_error(ans[:, :3, 3], gt[:, :3, 3])
        ms = L.rmse_loss(src, ans, gt)
        side[name] = dict(sel=sel, best=m[idx, sel], R=R[idx, sel], T=T[idx, sel])
        rep[name] = dict(rot_err_deg_mean=float(re.mean()), rot_err_deg_median=float(re.median()),
                         trans_err_mean=float(te.mean()), trans_err_median=float(te.median()),
                         mse_mean=float(ms.mean()), mse_median=float(ms.median()),
                         best_min_1_mean=float(m[idx, sel].mean()))
    a, b = side["fused"], side["reference_loop"]
    rel = ((a["best"] - b["best"]).abs() / b["best"]).cpu().numpy()
    cosv = ((torch.einsum("bij,bij->b", a["R"], b["R"]) - 1) / 2).clamp(-1, 1)
    rep["per_pair"] = dict(
        same_selected_start=float((a["sel"] == b["sel"]).float().mean()),
        best_min_1_rel_median=float(np.median(rel)), best_min_1_rel_max=float(rel.max()),
        selected_rot_delta_deg_median=float(torch.rad2deg(torch.acos(cosv)).median()),
        selected_trans_delta_median=float((a["T"] - b["T"]).norm(dim=1).median()),
        # all 64 starts: how many end at the same min_1 (1e-3 relative)
        starts_same_end_1e_3=float((((m_our - m_ref).abs() / m_ref) <= 1e-3).float().mean()))
    sel2 = m_ref2.argmin(1)
    rel2 = ((m_ref2[idx, sel2] - b["best"]).abs() / b["best"]).cpu().numpy()
    rep["reference_loop_against_itself"] = dict(
        same_selected_start=float((sel2 == b["sel"]).float().mean()),
        best_min_1_rel_median=float(np.median(rel2)), best_min_1_rel_max=float(rel2.max()),
        starts_same_end_1e_3=float((((m_ref2 - m_ref).abs() / m_ref) <= 1e-3).float().mean()))
    return rep


def test_distribution_against_reference_gpu_loop():
    from oracle import ref_ext
    if not ref_ext.available():
        pytest.skip("oracle/_ref/chamfer_3D.so (the reference extension) was not built")
    rep = distribution_report()
    print("\nDISTRIBUTIONAL PARITY\n" + json.dumps(rep, indent=1))
    pp, f, r = rep["per_pair"], rep["fused"], rep["reference_loop"]
    # per pair: the selected start ends at the same quality (200 chaotic iterations: which of two
    # near-equal starts wins is not reproducible by the reference either)
    assert pp["best_min_1_rel_median"] <= 2e-2
    assert pp["same_selected_start"] >= 0.3
    assert pp["starts_same_end_1e_3"] >= 0.5
    # over the set: the metrics the reference reports (train_utils.py:82-95)
    assert abs(f["best_min_1_mean"] - r["best_min_1_mean"]) <= 0.02 * r["best_min_1_mean"]
    assert abs(f["rot_err_deg_mean"] - r["rot_err_deg_mean"]) <= max(0.02 * r["rot_err_deg_mean"], 0.05)
    assert abs(f["trans_err_mean"] - r["trans_err_mean"]) <= max(0.02 * r["trans_err_mean"], 5e-4)
    assert abs(f["mse_mean"] - r["mse_mean"]) <= max(0.02 * r["mse_mean"], 5e-4)
    assert abs(f["rot_err_deg_median"] - r["rot_err_deg_median"]) <= 0.05
    assert abs(f["trans_err_median"] - r["trans_err_median"]) <= 5e-4
    assert abs(f["mse_median"] - r["mse_median"]) <= 5e-4
