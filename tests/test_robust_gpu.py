"""Robust noise models (noiseModel::Robust with the m-estimators of linear/LossFunctions.cpp) against the unmodified
reference: errors, whitened and reweighted Jacobians, one damped solve and a whole LM run per case.  SURVEY 8f row 3."""
import numpy as np
import pytest

import make_robust_golden
from conftest import load_golden
from gtsam_petercdev_b200.engine import Engine
from gtsam_petercdev_b200.problem import BUCKETS

pytestmark = pytest.mark.gpu

_problems = {}


def problem(name):
    if name not in _problems:
        import make_golden
        p = make_robust_golden.make_problem(name)
        g = load_golden("robust__" + name)
        assert make_golden.problem_digest(p) == str(g["digest"])
        _problems[name] = p
    return _problems[name]


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


@pytest.mark.parametrize("name", sorted(make_robust_golden.CASES))
def test_robust_linearize_and_solve(name):
    g = load_golden("robust__" + name)
    e = Engine(problem(name), device=0)
    assert abs(e.error(0) - g["lin_error0"][0]) <= 1e-12 * g["lin_error0"][0]        # sum of rho(|whitened residual|)
    e.linearize()
    for b in BUCKETS:
        if "lin_" + b + "_Ab" in g:
            ours, ref = e.jacobians(b), g["lin_" + b + "_Ab"]
            scale = np.abs(ref).max(axis=1, keepdims=True) + 1e-300
            # rows whose weight is exactly zero (Tukey, dead zone) are all-zero on both sides
            assert np.max(np.abs(ours - ref) / np.maximum(scale, 1e-200)) < 1e-11, b
    if int(g["solve_failed"][0]):
        return
    e.set_ordering(g["solve_ordering"])
    e.factorize(1e-3, False)
    e.backsolve()
    # bundle adjustment at lambda = 1e-3 without diagonal scaling is ill-conditioned
    # (the dead-zone case has weight 0 on every projection: the solution is ~1e-18, hence the absolute floor)
    tol = 1e-9 if not name.startswith("ba_") else 1e-6
    assert np.linalg.norm(e.delta() - g["solve_delta"]) <= tol * np.linalg.norm(g["solve_delta"]) + 1e-12
    e0, e1 = e.linear_errors()
    assert abs(e0 - g["solve_linear_error0"][0]) <= 1e-11 * g["solve_linear_error0"][0] + 1e-12
    assert abs(e1 - g["solve_linear_error1"][0]) <= 1e-7 * g["solve_linear_error0"][0] + 1e-12


@pytest.mark.parametrize("name", sorted(make_robust_golden.CASES))
def test_robust_lm_run(name):
    g = load_golden("robust__" + name)
    e = Engine(problem(name), device=0)
    e.set_ordering(g["lm_ordering"])
    r = e.lm_optimize()
    assert r["iterations"] == int(g["lm_iterations"][0])
    assert r["inner_iterations"] == int(g["lm_inner_iterations"][0])
    assert abs(r["final_error"] - g["lm_final_error"][0]) <= 1e-6 * g["lm_final_error"][0]


def test_robust_parameter_is_validated():
    from gtsam_petercdev_b200.engine import EngineError
    e = Engine(problem("pose2_300_huber"), device=0)
    with pytest.raises(EngineError):
        e.set_robust("between2", "huber", 0.0)       # the reference's constructors throw for k <= 0
    with pytest.raises(EngineError):
        e.set_robust("between2", 99, 1.0)
    e.set_robust("between2", "none", 0.0)
    g = load_golden("pose2_300__linearize")
    assert abs(e.error(0) - g["error0"][0]) <= 1e-12 * g["error0"][0]     # back to the plain Gaussian error
