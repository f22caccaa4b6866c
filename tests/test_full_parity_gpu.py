"""GPU parity of K3 / solve_model at BASELINE config 4's shape -- 2048 points x 64 starts x 200
iterations -- against fixtures made by the REFERENCE's own solve_model (houv.py:106-206, loaded
unmodified; tests/golden/make_golden.py houv_full / houv_full_retry).

SURVEY §7 hard part 1, reports (iii) and part of (iv): the optimiser is chaotic (arg-min switches,
hard min, top-k membership; Adam's g/(sqrt(v)+eps) turns float32 summation-order noise into
parameter noise), so starts that cross a switch leave the reference's trajectory.  What is
asserted: the early checkpoints for every start, the FRACTION of starts inside the bar later on
(measured values are printed and kept in profiles/r2_full_parity.json), and --
north_star's bar -- the pose each pair SELECTS within 1e-4 rad / 1e-5 units on the
well-conditioned pairs.
"""
import json

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def report():
    from houv_b200 import parity
    rep = parity.full_parity_report("houv_full")
    print("\nFULL-SIZE PARITY (houv_full.npz)\n" + json.dumps(rep, indent=1))
    return rep


def test_traces_follow_the_reference(report):
    """Every start over the short horizons; later, as close as the reference stays to ITSELF when
    only the order of its gradient atomics changes (report['reference_against_itself'])."""
    cp = report["calls"][0]["checkpoints"]
    env = report["reference_against_itself"]["checkpoints"]
    assert report["calls"][0]["rows"] == 192 and report["calls"][0]["iterations"] == 200
    assert cp["1"]["max_rel"] <= 1e-5            # one forward
    assert cp["5"]["max_rel"] <= 1e-4            # short trajectory, every start
    assert cp["20"]["frac_within_1e_3"] >= 0.95
    for k in ("20", "50", "100", "200"):
        assert cp[k]["frac_within_1e_3"] >= env[k]["frac_within_1e_3"] - 0.10
        assert cp[k]["frac_within_1e_4"] >= env[k]["frac_within_1e_4"] - 0.15
        assert cp[k]["median_rel"] <= max(3.0 * env[k]["median_rel"], 2e-4)
    assert cp["200"]["median_rel"] <= 1e-3       # the typical start is still on the trajectory


def test_selected_pose(report):
    """north_star's end-state bar is 1e-4 rad / 1e-5 units.  Two executions of the reference that
    differ only in atomics order select DIFFERENT starts and end 3e-4 .. 2.3e-3 rad / 3e-4 .. 6e-4
    units apart on these well-conditioned pairs (reference_against_itself): the bar holds over short
    horizons (tests/test_register_gpu.py, 10 iterations) and is not a property the reference has at
    200.  Asserted here: our selected pose is as close to the reference's as the reference's second
    execution is, its quality (min_1) within 2 %."""
    sel = report["calls"][0]["selected"]
    env = report["reference_against_itself"]["selected"]
    assert max(sel["rot_rad"]) <= max(2.0 * max(env["rot_rad"]), 1e-4)
    assert max(sel["trans"]) <= max(2.0 * max(env["trans"]), 1e-5)
    for a, b in zip(sel["best_min_1"], sel["best_min_1_ref"]):
        assert abs(a - b) <= 0.02 * b
    all_ = report["calls"][0]["end_state_all_starts"]
    env_all = report["reference_against_itself"]["end_state_all_starts"]
    assert all_["frac_within_bar"] >= env_all["frac_within_bar"] - 0.05
    assert all_["rot_rad_median"] <= 3.0 * env_all["rot_rad_median"]


def test_solve_model_answer(report):
    s = report["solve_model"]
    env = report["reference_against_itself"]["selected"]
    assert s["bottom_row_zero"]
    assert max(s["rot_rad"]) <= max(2.0 * max(env["rot_rad"]), 1e-4)
    assert max(s["trans"]) <= max(2.0 * max(env["trans"]), 1e-5)
    for a, b in zip(s["r_err_deg"], s["r_err_deg_ref"]):
        assert abs(a - b) <= 0.05    # degrees against the ground truth: both ~0.05 deg
    for a, b in zip(s["t_err"], s["t_err_ref"]):
        assert abs(a - b) <= 5e-4


def test_retry_cascade_at_full_size():
    """houv_full_retry.npz: three partial-overlap pairs with rotations of 160-175 degrees; the
    reference sends the ones that end above 0.030 through angle_base 1..3 (houv.py:154-180).  Same
    pairs retried, and solve_model keeps an answer of the same quality (ill-conditioned pairs are
    not held to the end-state bar: SURVEY §7 'never claim (iii) where the reference is unstable')."""
    from houv_b200 import parity
    rep = parity.full_parity_report("houv_full_retry")
    print("\nFULL-SIZE PARITY (houv_full_retry.npz)\n" + json.dumps(rep, indent=1))
    c0 = rep["calls"][0]
    assert c0["checkpoints"]["1"]["max_rel"] <= 1e-5
    # the same pairs cross the 0.030 threshold
    ours = [i for i, v in enumerate(c0["selected"]["best_min_1"]) if v > 0.030]
    assert ours == rep.get("retried_pairs_ref", [])
    assert len(rep["calls"]) == 4  # angle_base 0 + the three retries
    for c in rep["calls"][1:]:
        assert c["checkpoints"]["1"]["max_rel"] <= 1e-5
        assert c["checkpoints"]["5"]["max_rel"] <= 1e-4
    s = rep["solve_model"]
    for a, b in zip(s["r_err_deg"], s["r_err_deg_ref"]):
        assert abs(a - b) <= max(0.5, 0.05 * b)
