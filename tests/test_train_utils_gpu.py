"""GPU: the functional solver of registration/train_utils.py (getPredict_angle :359-456, solve
:467-572 -- the submission path of test.py:64) through houv_b200.train_utils, i.e. K3 with
float64 state, lr 0.1, |T| = sin(tran_s*pi) and the 3-D term only.  Checked against fixtures
made by the reference's own code (tests/golden/solve_ref.npz) and against the oracle."""
import numpy as np
import pytest
import torch

from houv_b200 import synth
from oracle import houv as oh
from tests.util import load_npz

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=[13, 0], ids=["k3-symmetric", "k3-two-scans"])
def pose_variant(request):
    from houv_b200 import _lib
    lib = _lib.load()
    lib.houv_set_pose_variant(request.param)
    yield request.param
    lib.houv_set_pose_variant(-1)


def _geodesic_rad(Ra, Rb):
    M = np.einsum("...ji,...jk->...ik", Ra.astype(np.float64), Rb.astype(np.float64))
    v = 0.5 * np.stack([M[..., 2, 1] - M[..., 1, 2], M[..., 0, 2] - M[..., 2, 0],
                        M[..., 1, 0] - M[..., 0, 1]], -1)
    return np.arctan2(np.linalg.norm(v, axis=-1), (np.trace(M, axis1=-2, axis2=-1) - 1) / 2)


def test_initial_variables_follow_numpy_stream():
    from houv_b200 import train_utils as tu
    np.random.seed(5)
    p = tu.initial_variables(7)
    np.random.seed(5)
    ref = [np.random.randn(7, 3), np.random.randn(7, 1), np.random.randn(7, 3), np.random.randn(7, 1)]
    np.random.randn(7, 3)
    after = np.random.randn(1)
    assert p.dtype == np.float64 and np.array_equal(p, np.concatenate(ref, 1))
    np.random.seed(5)
    tu.initial_variables(7)
    assert np.array_equal(np.random.randn(1), after)  # the unused angle_XYZ_c draw is consumed


@pytest.mark.parametrize("base,its", [(0, 1), (0, 4), (2, 4)])
def test_getPredict_angle_matches_reference_golden(base, its):
    from houv_b200 import train_utils as tu
    g = load_npz("solve_ref.npz")
    src, tgt = torch.from_numpy(g["src"]).cuda(), torch.from_numpy(g["tgt"]).cuda()
    np.random.seed(77)
    l, R, T, ts = tu.getPredict_angle(src, tgt, kernel=16, num_epochs=its, angle_base=base)
    k = "gpa_b%d_i%d_" % (base, its)
    assert l.shape == (3, 16) and R.shape == (3, 16, 3, 3) and T.shape == (3, 16, 3)
    assert ts.shape == (48, 1)
    ref = g[k + "min_1"]
    assert np.abs(l.cpu().numpy() - ref).max() < 2e-5 * np.abs(ref).max()
    assert _geodesic_rad(R.cpu().numpy(), g[k + "R"]).max() < 1e-4
    # lr is 0.1 here (10x HOUV's): Adam's g/(sqrt(v)+eps) turns float32 summation-order noise
    # in the gradients into parameter noise ~1e-6 per step, so the translation bar is 5e-5 for
    # the worst start and 1e-5 (north_star) for the typical one
    dT = np.abs(T.cpu().numpy() - g[k + "T"]).max(-1)
    assert dT.max() < 5e-5 and np.median(dT) < 1e-5, (dT.max(), np.median(dT))
    assert np.abs(ts.cpu().numpy() - g[k + "tran_s"]).max() < 5e-5


def test_solve_matches_reference_golden():
    """solve with the inner horizon shortened to 6 like the fixture; every pair is retried at
    angle_base 1..3.  Bars: 1e-4 rad (north_star) / 1e-4 units (lr 0.1, see above)."""
    from houv_b200 import train_utils as tu
    g = load_npz("solve_ref.npz")
    src, tgt, gt = (torch.from_numpy(g[k]).cuda() for k in ("src", "tgt", "transform"))
    old, tu.SOLVE_EPOCHS = tu.SOLVE_EPOCHS, 6
    try:
        np.random.seed(78)
        r_err, t_err, ans = tu.solve(src, tgt, gt, kernel=16)
        np.random.seed(78)
        cpu = tu.solve(src, tgt, None, kernel=16, prefix='test')
    finally:
        tu.SOLVE_EPOCHS = old
    a = ans.cpu().numpy()
    assert np.abs(a[:, 3]).max() == 0
    assert _geodesic_rad(a[:, :3, :3], g["solve_ans"][:, :3, :3]).max() < 1e-4
    assert np.abs(a[:, :3, 3] - g["solve_ans"][:, :3, 3]).max() < 1e-4  # 6 steps at lr 0.1
    assert np.abs(t_err.cpu().numpy() - g["solve_t_err"]).max() < 1e-4
    assert not cpu.is_cuda and torch.equal(cpu, ans.cpu())


def test_float64_state_trajectory_vs_oracle():
    """20 steps at lr 0.1 on a larger cloud: the loss trace follows the float64-leaf oracle."""
    from houv_b200 import solver, train_utils as tu
    s, t, _ = synth.make_pairs(11, 2, n_points=256)
    src, tgt = torch.from_numpy(s), torch.from_numpy(t)
    np.random.seed(3)
    trace = []
    oh.getPredict_angle(src, tgt, kernel=8, num_epochs=20, trace=trace)
    np.random.seed(3)
    params = torch.from_numpy(tu.initial_variables(16)).cuda()
    m = torch.zeros(16, 8, dtype=torch.float64).cuda()
    v = torch.zeros(16, 8, dtype=torch.float64).cuda()
    out = solver.pose_optimise(src.cuda(), tgt.cuda(), params, 8, 20, lr=0.1, wview=0.0,
                               tran_scale=1.0, tran_offset=0.0, adam_m=m, adam_v=v,
                               want_trace=True)
    got = out["trace"][:, :, 1].cpu().numpy()
    ref = np.stack([x["min_1"].numpy() for x in trace])
    assert np.abs(got[:5] - ref[:5]).max() < 1e-5 * np.abs(ref).max()
    assert np.abs(got - ref).max() < 5e-3 * np.abs(ref).max()
    # first-step gradients from the float64 first moment: m_1 = (1-beta1) g
    g_gpu = (m.cpu().numpy() / (1 - 0.9 ** 20))  # only its sign/size is sanity-checked below
    assert np.isfinite(g_gpu).all() and params.dtype == torch.float64


def test_one_metric_sweep_equals_four_metric_sweep():
    """wview == 0 selects the 3-D-only specialisation of the symmetric sweep; variant 15 forces the
    four-metric sweep on the same problem: identical bits (the view terms never reach the loss)."""
    from houv_b200 import _lib, solver, train_utils as tu
    s, t, _ = synth.make_pairs(21, 2, n_points=320)
    src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
    outs = []
    for variant in (12, 15, 13):
        _lib.load().houv_set_pose_variant(variant)
        np.random.seed(9)
        params = torch.from_numpy(tu.initial_variables(32)).cuda()
        o = solver.pose_optimise(src, tgt, params, 16, 12, lr=0.1, wview=0.0, tran_scale=1.0,
                                 tran_offset=0.0, want_trace=True, want_tran_s=True)
        outs.append((params.clone(), o))
    p0, o0 = outs[0]
    for p1, o1 in outs[1:]:  # 13 = the one-metric sweep with four rows evaluated before their folds
        assert torch.equal(p0, p1) and torch.equal(o0["trace"], o1["trace"])
        assert torch.equal(o0["R"], o1["R"]) and torch.equal(o0["tran_s"], o1["tran_s"])


def test_getPredict_angle_full_horizon_against_reference_golden(pose_variant):
    """solve_full.npz: the reference's own getPredict_angle at the shape test.py:64 runs it -- 2048
    points x 64 starts x 500 Adam iterations, lr 0.1, float64 leaves (tests/golden/make_golden.py
    solve_full).  lr is 10x HOUV's, so trajectories leave the reference's sooner than in
    tests/test_full_parity_gpu.py (Adam's g/(sqrt(v)+eps) amplifies float32 summation-order noise;
    the reference itself adds with float atomics in no fixed order).  Asserted: every start over the
    first forwards; at 500 the 64 starts of a pair as a POPULATION (percentiles of the final min_1)
    and each pair's best start.  Measured on a B200: forward 1 / 3 max 3.3e-7 / 3.2e-6; forward 10
    median 1.3e-5; from forward 50 on single trajectories are unrelated (median 4-16 %) -- the
    reference's own min_1 of one start moves by 7.8 % (median) between its last two forwards, lr 0.1
    keeps every start oscillating; final percentiles 10/25/50/75/90: 0.0114 0.0133 0.0174 0.0195
    0.0235 (ours) against 0.0110 0.0135 0.0176 0.0196 0.0263 (reference); best start 0.0092 / 0.0079
    against 0.0087 / 0.0089."""
    if pose_variant != 13:
        pytest.skip("one fused kernel is enough for the 500-iteration run (the one-metric sweep)")
    from houv_b200 import solver, train_utils as tu
    g = load_npz("solve_full.npz")
    src, tgt = torch.from_numpy(g["src"]).cuda(), torch.from_numpy(g["tgt"]).cuda()
    ref = g["trace"]                                   # [500, 128]
    its, rows = ref.shape
    np.random.seed(int(g["seed"]))
    params = torch.from_numpy(tu.initial_variables(rows)).cuda()
    out = solver.pose_optimise(src, tgt, params, 64, its, angle_base=0, lr=tu.ADAM_LR, wview=0.0,
                               tran_scale=1.0, tran_offset=0.0, want_tran_s=True, want_trace=True)
    tr = out["trace"][:, :, 1].double().cpu().numpy()
    rel = np.abs(tr - ref) / np.abs(ref)
    rep = {k: dict(max=float(rel[k - 1].max()), median=float(np.median(rel[k - 1])),
                   within_1e_3=float((rel[k - 1] <= 1e-3).mean())) for k in (1, 3, 10, 50, 200, 500)}
    m1, m1r = out["min_1"].reshape(2, 64).cpu().numpy(), g["min_1"]
    best, best_ref = m1.min(1), m1r.min(1)
    q = [10, 25, 50, 75, 90]
    quant = dict(ours=[float(x) for x in np.percentile(m1, q)], reference=[float(x) for x in np.percentile(m1r, q)])
    # how much min_1 of ONE start moves between consecutive forwards near the end (lr 0.1 keeps it oscillating)
    osc = float(np.median(np.abs(ref[-1] - ref[-2]) / ref[-1]))
    print("\nN1 FULL HORIZON", rep, "best min_1", best, "reference", best_ref, "final min_1 percentiles",
          q, quant, "reference's own step-to-step change at the end (median)", osc)
    assert rep[1]["max"] <= 1e-5 and rep[3]["max"] <= 1e-4
    assert rep[10]["median"] <= 1e-4
    # at 500 forwards the 64 starts of a pair are compared as a population, and the best start as the result
    for a, b in zip(quant["ours"], quant["reference"]):
        assert abs(a - b) <= 0.15 * b
    assert np.abs(best - best_ref).max() <= 0.15 * best_ref.max()
