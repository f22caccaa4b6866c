"""oracle/emd_oracle.c (CPU restatement of emd_cuda.cu:95-266) against ground truth it does not depend on:
the optimal assignment from scipy (Hungarian).  The reference ships no test vectors for EMD beyond
emd_module.py:95-108 `test_emd` (which needs its GPU); the GPU suite compares against the reference extension."""
import numpy as np
from scipy.optimize import linear_sum_assignment

from oracle import emd as oe


def test_auction_reaches_the_optimum_within_eps():
    rng = np.random.RandomState(0)
    x1 = rng.rand(3, 200, 3).astype(np.float32)
    x2 = rng.rand(3, 200, 3).astype(np.float32)
    d, a, used = oe.emd_forward(x1, x2, 0.0005, 20000)
    for b in range(3):
        assert used[b] < 20000 and len(set(a[b])) == 200
        C = np.sqrt(((x1[b][:, None] - x2[b][None]) ** 2).sum(-1))
        r, c = linear_sum_assignment(C)
        assert C[r, c].mean() - 1e-6 <= np.sqrt(d[b]).mean() <= C[r, c].mean() + 0.0005 + 1e-6
        assert np.allclose(d[b], ((x1[b] - x2[b][a[b]]) ** 2).sum(-1), atol=1e-6)


def test_last_iteration_assigns_everything_and_identity_case():
    rng = np.random.RandomState(1)
    x = rng.rand(1, 64, 3).astype(np.float32)
    d, a, used = oe.emd_forward(x, x, 0.005, 50)
    assert (a[0] == np.arange(64)).all() and (d == 0).all() and used[0] == 1
    y = rng.rand(1, 64, 3).astype(np.float32)
    d, a, used = oe.emd_forward(x, y, 0.005, 1)  # one iteration = forced assignment to the best bid
    assert (a >= 0).all() and used[0] == 1
    g = oe.emd_backward(x, y, np.ones((1, 64), np.float32), a)
    assert np.allclose(g, 2 * (x - y[0][a[0]][None]), atol=1e-6)
