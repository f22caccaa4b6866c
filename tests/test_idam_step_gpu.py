"""K6 (houv_idam_step) against reference-made golden vectors (tests/golden/idam_ref.npz: the reference's own
IDAM SVDHead, models/idam.py:150-189, driving the restated iteration tail :317-342) and against the oracle on
random inputs.  Tolerances: weights 1e-6 absolute (float32 sigmoid / sum order), R and t 2e-5 (float32 SVD in the
reference against our float64 Jacobi), moved cloud 5e-5."""
import numpy as np
import pytest
import torch

from oracle import houv as oh
from tests.util import load_npz


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["even", "odd"])
def test_three_chained_steps_match_reference_golden(tag):
    from houv_b200.kabsch import idam_step
    g = load_npz("idam_ref.npz")
    tgt = _dev(g[tag + "_tgt"])
    for it in range(3):
        k = lambda n: g["%s_%s%d" % (tag, n, it)]
        out, R, t, Rab, tab, w = idam_step(_dev(k("src")), tgt, _dev(k("corr")), _dev(k("logits")),
                                           _dev(k("Rin")), _dev(k("tin")))
        assert np.abs(w.cpu().numpy() - k("w")).max() < 1e-6
        assert ((w.cpu().numpy() > 0) == (k("w") > 0)).all()  # same points eliminated
        assert np.abs(Rab.cpu().numpy() - k("Rab")).max() < 2e-5
        assert np.abs(tab.cpu().numpy() - k("tab")).max() < 2e-5
        assert np.abs(R.cpu().numpy() - k("R")).max() < 2e-5
        assert np.abs(t.cpu().numpy() - k("t")).max() < 2e-5
        assert np.abs(out.cpu().numpy() - k("out")).max() < 5e-5


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 1, 4), (2, 2, 3), (7, 33, 50), (300, 128, 128), (4, 2048, 1500)])
@pytest.mark.parametrize("idx_dtype", [torch.int64, torch.int32])
def test_random_steps_match_oracle(shape, idx_dtype):
    from houv_b200.kabsch import idam_step
    B, N, M = shape
    torch.manual_seed(N * 7 + M)
    src = torch.randn(B, 3, N) * 0.3
    tgt = torch.randn(B, 3, M) * 0.3
    corr = torch.randint(0, M, (B, N))
    logits = torch.randn(B, N) * 2
    R0 = torch.linalg.qr(torch.randn(B, 3, 3))[0]
    t0 = torch.randn(B, 3)
    exp = oh.idam_step(src, tgt, corr, logits, R0, t0)
    got = idam_step(src.cuda(), tgt.cuda(), corr.to(idx_dtype).cuda(), logits.cuda(), R0.cuda(), t0.cuda())
    ew, gw = exp[5].numpy(), got[5].cpu().numpy()
    assert np.abs(gw - ew).max() < 1e-6 and ((gw > 0) == (ew > 0)).all()
    if N >= 8:  # tiny clouds are rank-deficient: R is not unique there, only the weights are checked
        for e, g_, tol in zip(exp[:5], got[:5], (1e-4, 5e-5, 5e-5, 5e-5, 5e-5)):
            assert np.abs(g_.cpu().numpy() - e.numpy()).max() < tol
    for g_ in got:
        assert torch.isfinite(g_).all()


@pytest.mark.gpu
def test_defaults_identity_pose_and_bad_inputs():
    from houv_b200.kabsch import idam_step, hybrid_point_elimination
    torch.manual_seed(3)
    src, tgt = torch.randn(2, 3, 64).cuda(), torch.randn(2, 3, 64).cuda()
    corr = torch.randint(0, 64, (2, 64)).cuda()
    logits = torch.randn(2, 64).cuda()
    out, R, t, Rab, tab, w = idam_step(src, tgt, corr, logits)
    assert torch.allclose(R, Rab) and torch.allclose(t, tab)  # identity start (idam.py:262-263)
    assert torch.allclose(w, hybrid_point_elimination(logits), atol=1e-6)
    assert torch.allclose(w.sum(-1), torch.ones(2).cuda(), atol=1e-5)
    with pytest.raises(RuntimeError):
        idam_step(src.cpu(), tgt, corr, logits)
    with pytest.raises(RuntimeError):
        idam_step(src, tgt, corr.float(), logits)
    with pytest.raises(RuntimeError):
        idam_step(src, tgt, corr, logits[:, :10])
    with pytest.raises(RuntimeError):  # N above the shared-memory cap -> HOUV_E_TOOBIG
        idam_step(torch.randn(1, 3, 9000).cuda(), tgt[:1], torch.zeros(1, 9000, dtype=torch.int64).cuda(),
                  torch.zeros(1, 9000).cuda())
