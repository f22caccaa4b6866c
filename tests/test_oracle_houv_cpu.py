"""CPU: pins oracle/houv.py (restated registration/models/houv.py & co.) against golden
vectors produced by the reference's OWN python files run unmodified in the dev container
(tests/golden/make_golden.py houv|svd through tests/golden/ref_loader.py).

Same torch ops in the same order, so the values agree to float rounding of the CPU's
vector width (tolerance 1e-6 relative); discrete outputs (initial draws) are exact."""
import numpy as np
import torch

from oracle import houv as oh
from tests.util import load_npz

RTOL = 2e-6


def _close(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() <= rtol * max(np.abs(b).max(), 1e-30)


def test_reset_weight_draws_exact():
    g = load_npz("houv_ref.npz")
    vc, ac, tc, ts = oh.initial_parameters(64)
    assert np.array_equal(vc, g["init_V_c"]) and np.array_equal(ac, g["init_angle_c"])
    assert np.array_equal(tc, g["init_tran_c"]) and np.array_equal(ts, g["init_tran_s"])
    assert np.array_equal(ac, ts)  # numpy re-seeded before each tensor (houv.py:54-60)
    assert np.array_equal(vc[0], [-1, -1, -1]) and np.array_equal(vc[25], [1, 1, 1])


def test_forward_and_predict_loss():
    g = load_npz("houv_ref.npz")
    src, tgt = torch.from_numpy(g["src"]), torch.from_numpy(g["tgt"])
    rep = lambda x: x.unsqueeze(1).expand(-1, 32, -1, -1).reshape(-1, 128, 3)
    net = oh.HOUV()
    net.reset_weight(64, 0)
    src_temp, R, T = net(rep(src))
    loss, min_1 = oh.Predict_loss(src_temp, rep(tgt))
    assert _close(src_temp.detach(), g["fwd_src_temp"]) and _close(R.detach(), g["fwd_R"])
    assert _close(T.detach(), g["fwd_T"])
    assert _close(loss.detach(), g["fwd_loss"]) and _close(min_1.detach(), g["fwd_min_1"])


def test_predict_model_three_iterations():
    g = load_npz("houv_ref.npz")
    src, tgt = torch.from_numpy(g["src"]), torch.from_numpy(g["tgt"])
    for base in (0, 2):
        l, R, T = oh.predict_model(oh.HOUV(), src, tgt, kernel=32, num_epochs=3, angle_base=base)
        assert _close(l.detach(), g["pm%d_loss" % base], 1e-5)
        assert _close(R.detach(), g["pm%d_R" % base], 1e-5)
        assert _close(T.detach(), g["pm%d_T" % base], 1e-5)


def test_solve_model_and_metrics():
    g = load_npz("houv_ref.npz")
    src, tgt, gt = (torch.from_numpy(g[k]) for k in ("src", "tgt", "transform"))
    r_err, t_err, ans = oh.solve_model(oh.HOUV(), src, tgt, gt, kernel=32, num_epochs=10)
    assert _close(ans, g["solve_ans"], 1e-4)
    assert np.abs(ans[:, 3].numpy()).max() == 0  # zero bottom row, houv.py:187-195
    assert _close(r_err, g["solve_r_err"], 1e-3) and _close(t_err, g["solve_t_err"], 1e-4)
    assert _close(oh.rmse_loss(src, ans, gt), g["solve_mse"], 1e-4)


def test_svdhead():
    g = load_npz("svdhead_ref.npz")
    src, corr, w = (torch.from_numpy(g[k]) for k in ("src", "corr", "weights"))
    R, t = oh.SVDHead()(src, corr)
    assert _close(R, g["R"], 1e-5) and _close(t, g["t"], 1e-5)
    Rw, tw = oh.SVDHead()(src, corr, w)
    assert _close(Rw, g["Rw"], 1e-5) and _close(tw, g["tw"], 1e-5)
    assert torch.allclose(torch.det(R), torch.ones(6), atol=1e-5)  # reflection fixed


def test_icp_oracle_recovers_small_motion():
    """Unpinned restatement (no Open3D here): sanity only -- a 3 degree / 5 mm perturbation of a
    cloud against itself is recovered."""
    rng = np.random.RandomState(0)
    tgt = rng.rand(400, 3) - 0.5
    a = np.deg2rad(3.0)
    Rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    src = (tgt - np.array([0.005, 0, 0])) @ Rz  # tgt = Rz src + t
    T, fit, rm, it = oh.icp_point_to_point(src, tgt, max_distance=0.05, max_iteration=100)
    assert fit == 1.0 and rm < 1e-6 and it < 100
    assert np.abs(T[:3, :3] - Rz).max() < 1e-6


# ---------------------------------------------------------------- train_utils.py solver
def test_getPredict_angle_matches_reference_golden():
    """oracle getPredict_angle vs the reference's own (train_utils.py:359-456), NumPy seeded the
    same way: float64 leaves, lr 0.1, loss 6*min_1, |T| = sin(tran_s*pi)."""
    g = load_npz("solve_ref.npz")
    src, tgt = torch.from_numpy(g["src"]), torch.from_numpy(g["tgt"])
    for base, its in ((0, 1), (0, 4), (2, 4)):
        np.random.seed(77)
        l, R, T, ts = oh.getPredict_angle(src, tgt, kernel=16, num_epochs=its, angle_base=base)
        k = "gpa_b%d_i%d_" % (base, its)
        assert _close(l.detach(), g[k + "min_1"], 1e-5) and _close(R.detach(), g[k + "R"], 1e-5)
        assert _close(T.detach(), g[k + "T"], 1e-5) and _close(ts.detach(), g[k + "tran_s"], 1e-5)


def test_solve_matches_reference_golden():
    """oracle solve vs the reference's own solve (train_utils.py:467-572) with its inner
    500-iteration calls shortened to 6: all three pairs stay above 0.030 and go through the
    angle_base 1..3 retries, each drawing fresh values from NumPy's global generator."""
    g = load_npz("solve_ref.npz")
    src, tgt, gt = (torch.from_numpy(g[k]) for k in ("src", "tgt", "transform"))
    np.random.seed(78)
    r_err, t_err, ans = oh.solve(src, tgt, gt, kernel=16, epochs=6)
    assert _close(ans, g["solve_ans"], 1e-4)
    assert np.abs(ans[:, 3].numpy()).max() == 0
    assert _close(r_err, g["solve_r_err"], 1e-3) and _close(t_err, g["solve_t_err"], 1e-4)


def test_calc_cd_with_fscore_matches_reference_golden():
    """oracle calc_cd / fscore vs the reference's own (model_utils_completion.py:69-80,
    fscore.py:3-16), values and the gradient of mean(cd_p)+mean(cd_t) w.r.t. the output cloud."""
    g = load_npz("calccd_ref.npz")
    o = torch.from_numpy(g["output"]).requires_grad_()
    cd_p, cd_t, f1 = oh.calc_cd(o, torch.from_numpy(g["gt"]), calc_f1=True)
    (cd_p.mean() + cd_t.mean()).backward()
    assert _close(cd_p.detach(), g["cd_p"]) and _close(cd_t.detach(), g["cd_t"])
    assert np.array_equal(f1.numpy(), g["f1"])
    # coincident points (d = 0) make sqrt's backward 0 * inf = NaN in the reference: same places
    got, ref = o.grad.numpy(), g["grad_output"]
    assert np.isnan(ref).any() and np.array_equal(np.isnan(got), np.isnan(ref))
    assert _close(np.nan_to_num(got), np.nan_to_num(ref), 1e-5)


def test_idam_step_restatement_matches_reference_golden():
    """oracle.idam_step with the oracle's own SVDHead reproduces the vectors made with the reference's IDAM
    SVDHead class (tests/golden/make_golden.py idam)."""
    g = load_npz("idam_ref.npz")
    for tag in ("even", "odd"):
        tgt = torch.from_numpy(g[tag + "_tgt"])
        for it in range(3):
            k = lambda n: torch.from_numpy(g["%s_%s%d" % (tag, n, it)])
            got = oh.idam_step(k("src"), tgt, k("corr"), k("logits"), k("Rin"), k("tin"))
            for name, v in zip(("out", "R", "t", "Rab", "tab", "w"), got):
                assert torch.allclose(v, k(name), atol=1e-6), (tag, it, name)


def test_oracle_follows_reference_at_full_size():
    """houv_full.npz: the reference's own predict_model at BASELINE config 4's shape (2048 points x
    64 starts, 3 pairs).  The restated oracle must follow its min_1 / loss traces over the first
    iterations at the same size (the whole 200-iteration horizon is the GPU test's job)."""
    from oracle import chamfer as oc
    import os
    oc.set_threads(os.cpu_count() or 8)
    g = load_npz("houv_full.npz")
    src, tgt = torch.from_numpy(g["src"]), torch.from_numpy(g["tgt"])
    trace = []
    oh.predict_model(oh.HOUV(), src, tgt, kernel=64, num_epochs=4, angle_base=0, trace=trace)
    for k in range(4):
        assert _close(trace[k]["min_1"], g["pm0_trace"][k], 1e-5)
        assert _close(trace[k]["loss"], g["pm0_loss"][k], 1e-5)
