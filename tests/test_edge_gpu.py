"""Edge cases the reference's own unit tests pin, driven through the device path and compared with the unmodified
reference's dumps (tests/golden/make_edge_golden.py): the three SO3::Logmap branches (near pi, ordinary, near zero;
geometry/SO3.cpp:299-375, gtsam/geometry/tests/testSO3.cpp, testPose3.cpp), Pose3 Logmap's small-angle shortcut and Expmap's
nearZero branch (geometry/Pose3.cpp:189, 230-233), cheirality zero rows (slam/GeneralSFMFactor.h:127-158), and the matrices
of gtsam/base/tests/testCholesky.cpp:101-138 that must fail."""
import numpy as np
import pytest

import make_edge_golden as edge
import make_golden
from conftest import check_delta, load_golden
from gtsam_petercdev_b200.engine import Engine, IndeterminantLinearSystemException

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


def _problem(builder, tag, *args):
    p = builder(*args)
    assert make_golden.problem_digest(p) == str(load_golden(tag)["digest"]), "edge-case generator drifted from its fixture"
    return p


def _check_jacobians(e, g, buckets):
    assert abs(e.error(0) - g["error0"][0]) <= 1e-12 * g["error0"][0]
    e.linearize()
    for b in buckets:
        ours, ref = e.jacobians(b), g[b + "_Ab"]
        assert ours.shape == ref.shape
        scale = np.abs(ref).max(axis=1, keepdims=True) + 1e-300
        assert np.max(np.abs(ours - ref) / scale) < 1e-12, b
    return e


def _check_solve(e, g, tag, err_tol=1e-9):
    e.set_ordering(g["ordering"])
    assert np.array_equal(e.plan()["clique_keys"], g["clique_keys"])
    e.linearize()
    e.factorize(float(g["lam"]), False)
    e.backsolve()
    d = e.delta()
    check_delta(tag, d, g["delta"], g["delta_star"])
    e0, e1 = e.linear_errors()
    assert abs(e0 - g["linear_error0"][0]) <= 1e-12 * g["linear_error0"][0]
    assert abs(e1 - g["linear_error1"][0]) <= 1e-8 * g["linear_error0"][0]
    e.retract()
    assert abs(e.error(1) - g["error1"][0]) <= err_tol * max(g["error1"][0], g["error0"][0] * 1e-3)


def test_so3_logmap_branches_near_pi_and_near_zero():
    g = load_golden("edge_pose3_logmap__linearize")
    p = _problem(edge.pose3_logmap_problem, "edge_pose3_logmap__linearize")
    e = _check_jacobians(Engine(p, device=0), g, ["between3", "prior3"])
    # the fixture really covers the branches: residual rotation norms at pi - 1e-4, tiny, and zero
    b = g["between3_Ab"].reshape(-1, 6, 13)[:, :3, 12] * 0.05   # un-whitened rotation residual (sigma 0.05)
    ang = np.linalg.norm(b, axis=1)
    assert np.sum(np.abs(ang - (np.pi - 1e-4)) < 1e-7) >= 5 and np.sum((ang > 0) & (ang < 1e-6)) >= 2 and np.sum(ang < 1e-12) >= 1
    _check_solve(e, load_golden("edge_pose3_logmap__solve"), "edge_pose3_logmap__solve")


def test_pose3_expmap_near_zero_through_an_lm_run():
    g = load_golden("edge_pose3_logmap__lm")
    p = _problem(edge.pose3_logmap_problem, "edge_pose3_logmap__lm")
    e = Engine(p, device=0)
    e.set_ordering(g["ordering"])
    r = e.lm_optimize()
    assert r["iterations"] == int(g["iterations"][0]) and r["inner_iterations"] == int(g["inner_iterations"][0])
    assert abs(r["final_error"] - g["final_error"][0]) <= 1e-6 * g["final_error"][0]
    order = np.argsort(p.a["pose3_keys"], kind="stable")
    assert np.allclose(e.variables(1)[order], g["final_pose3"], rtol=0, atol=1e-6 * max(1.0, np.abs(g["final_pose3"]).max()))


def test_cheirality_gives_zero_rows_and_the_factor_drops_out():
    g = load_golden("edge_ba_behind__linearize")
    p = _problem(edge.ba_behind_problem, "edge_ba_behind__linearize")
    e = _check_jacobians(Engine(p, device=0), g, ["sfm", "priorcam", "priorpt"])
    ours, ref = e.jacobians("sfm"), g["sfm_Ab"]
    dropped = np.all(ref == 0.0, axis=1)
    assert dropped.sum() >= 10
    assert np.array_equal(np.all(ours == 0.0, axis=1), dropped)       # exactly the same factors, exactly zero
    # the error after the step (6e12: a few points end up next to a camera's principal plane) magnifies the 6e-10 difference
    # of the two deltas forty-fold: 1e-6 on that one scalar, the usual bars on everything else
    _check_solve(e, load_golden("edge_ba_behind__solve"), "edge_ba_behind__solve", err_tol=1e-6)


@pytest.mark.parametrize("which", ["A1", "A2"])
def test_matrices_that_must_fail(which):
    """gtsam/base/tests/testCholesky.cpp:101-138: choleskyPartial returns false on A1 (last pivot 1e-12: exponent test,
    base/cholesky.cpp:146-158) and A2 (rank deficient)."""
    g = load_golden(f"edge_chol_fail_{which}__solve")
    assert int(g["failed"][0]) == 1
    p = _problem(edge.chol_fail_problem, f"edge_chol_fail_{which}__solve", which)
    e = Engine(p, device=0)
    e.set_ordering("natural")
    e.linearize()
    with pytest.raises(IndeterminantLinearSystemException) as ei:
        e.factorize(0.0)
    assert ei.value.key == int(g["failed_key"][0])
