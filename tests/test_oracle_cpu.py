"""CPU: pins the oracle (oracle/chamfer_oracle.c) against the reference.

  * golden vectors produced by the reference's own code in the dev container
    (tests/golden/make_golden.py): chamfer_python.distChamfer and the reference
    CUDA kernel source built for the host;
  * the prebuilt oracle/_ref library itself when it is present.
Bars: idx bit-exact (utils/metrics/CD/unit_test.py:29-33), dist bit-exact vs the
kernel source and MSE < 1e-8 vs the float64 python form (:25-27).
"""
import numpy as np
import pytest

from oracle import chamfer as oc
from tests.util import grouped, load_npz


def test_unit_case_matches_reference_python_and_kernel():
    g = load_npz("chamfer_unit.npz")
    d1, d2, i1, i2 = oc.chamfer_forward(g["xyz1"], g["xyz2"])
    # the reference's own acceptance test
    mse = np.mean((d1 - g["py_dist1"]) ** 2) + np.mean((d2 - g["py_dist2"]) ** 2)
    assert mse < 1e-8
    assert np.array_equal(i1, g["py_idx1"]) and np.array_equal(i2, g["py_idx2"])
    # bit-exact against the kernel source
    assert np.array_equal(d1, g["ref_dist1"]) and np.array_equal(d2, g["ref_dist2"])
    assert np.array_equal(i1, g["ref_idx1"]) and np.array_equal(i2, g["ref_idx2"])
    g1, g2 = oc.chamfer_backward(g["xyz1"], g["xyz2"], g["graddist1"], g["graddist2"], i1, i2)
    assert np.array_equal(g1, g["ref_gradxyz1"]) and np.array_equal(g2, g["ref_gradxyz2"])


@pytest.mark.parametrize("case", sorted(grouped(load_npz("chamfer_cases.npz"))))
def test_edge_cases_match_reference_kernel(case):
    c = grouped(load_npz("chamfer_cases.npz"))[case]
    d1, d2, i1, i2 = oc.chamfer_forward(c["xyz1"], c["xyz2"])
    assert np.array_equal(i1, c["idx1"]) and np.array_equal(i2, c["idx2"])
    assert np.array_equal(d1, c["dist1"]) and np.array_equal(d2, c["dist2"])
    g1, g2 = oc.chamfer_backward(c["xyz1"], c["xyz2"], c["gd1"], c["gd2"], i1, i2)
    assert np.array_equal(g1, c["g1"]) and np.array_equal(g2, c["g2"])
    h1, h2 = oc.chamfer_backward(c["xyz1"], c["xyz2"], c["gd1"], c["gd2"], i1, i2, f64=True)
    scale = max(np.abs(h1).max(), np.abs(h2).max(), 1e-30)
    assert np.abs(g1 - h1).max() <= 1e-5 * scale and np.abs(g2 - h2).max() <= 1e-5 * scale


def test_tie_break_is_first_minimum():
    c = grouped(load_npz("chamfer_cases.npz"))["dup_targets"]
    # queries 0,1 sit next to targets 5 and 510, which are duplicated later on
    assert c["idx1"][0, 0] == 5 and c["idx1"][0, 1] == 510


@pytest.mark.skipif(not oc.ref_available(), reason="oracle/_ref not built")
def test_against_prebuilt_reference_kernels_random():
    rng = np.random.RandomState(3)
    for (b, n, m) in [(2, 64, 700), (1, 1025, 1024), (3, 5, 3)]:
        x1 = rng.rand(b, n, 3).astype(np.float32)
        x2 = rng.rand(b, m, 3).astype(np.float32)
        a, r = oc.chamfer_forward(x1, x2), oc.ref_chamfer_forward(x1, x2)
        for u, v in zip(a, r):
            assert np.array_equal(u, v)
        gd1 = rng.randn(b, n).astype(np.float32)
        gd2 = rng.randn(b, m).astype(np.float32)
        ga = oc.chamfer_backward(x1, x2, gd1, gd2, a[2], a[3])
        gr = oc.ref_chamfer_backward(x1, x2, gd1, gd2, a[2], a[3])
        for u, v in zip(ga, gr):
            assert np.array_equal(u, v)


def test_threads_do_not_change_results():
    rng = np.random.RandomState(5)
    x1 = rng.rand(3, 301, 3).astype(np.float32)
    x2 = rng.rand(3, 222, 3).astype(np.float32)
    oc.set_threads(1)
    a = oc.chamfer_forward(x1, x2)
    oc.set_threads(4)
    b = oc.chamfer_forward(x1, x2)
    oc.set_threads(1)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)


def test_nan_semantics_follow_the_reference_kernel():
    """chamfer3D.cu:36,79,121,126: a NaN distance enters `best` only at position 0 of a 512-point
    tile (k == 0 || d < best) and then blocks that whole tile; across tiles `result > best` skips a
    NaN tile but can never leave a NaN first tile.  The C restatement (scalar and AVX2 paths) must
    reproduce the reference kernel source on NaNs at tile positions 0 / 5 / 512 / 513, a NaN
    query, and +-inf coordinates.  (The CUDA kernels take finite inputs only: include/houv_b200.h.)"""
    if not oc.ref_available():
        import pytest
        pytest.skip("oracle/_ref not built")
    rng = np.random.RandomState(5)
    base1 = rng.rand(1, 40, 3).astype(np.float32)
    base2 = rng.rand(1, 1100, 3).astype(np.float32)
    cases = []
    for pos in (0, 5, 512, 513, 1024):
        b = base2.copy()
        b[0, pos, 1] = np.nan
        cases.append((base1, b))
    a = base1.copy()
    a[0, 7, 0] = np.nan
    cases.append((a, base2))
    b = base2.copy()
    b[0, 512] = [np.inf, 0, 0]
    b[0, 3] = [-np.inf, np.inf, 0]
    cases.append((base1, b))
    for simd in (1, 0):
        oc.oracle_lib().oracle_set_simd(simd)
        for x1, x2 in cases:
            got, ref = oc.chamfer_forward(x1, x2), oc.ref_chamfer_forward(x1, x2)
            for g, r in zip(got, ref):
                assert np.array_equal(g, r, equal_nan=True)
    oc.oracle_lib().oracle_set_simd(1)
    # what the semantics are, spelled out once: NaN at tile position 0 of the FIRST tile poisons
    # every query; at position 512 it hides the whole second tile; at 5 or 513 only that point
    b = base2.copy(); b[0, 0, 1] = np.nan
    d1, _, i1, _ = oc.ref_chamfer_forward(base1, b)
    assert np.isnan(d1).all() and (i1 == 0).all()
    b = base2.copy(); b[0, 512, 1] = np.nan
    _, _, i1, _ = oc.ref_chamfer_forward(base1, b)
    assert not ((i1 >= 512) & (i1 < 1024)).any()


def test_simd_and_scalar_paths_agree():
    rng = np.random.RandomState(6)
    x1 = rng.rand(3, 203, 3).astype(np.float32)
    x2 = (rng.randint(0, 5, (3, 517, 3)) / 4.0).astype(np.float32)  # many ties
    oc.oracle_lib().oracle_set_simd(0)
    a = oc.chamfer_forward(x1, x2)
    oc.oracle_lib().oracle_set_simd(1)
    b = oc.chamfer_forward(x1, x2)
    for p, q in zip(a, b):
        assert np.array_equal(p, q)
