"""The hot path at BASELINE.json's full sizes (C1 M3500-shape, C2 sphere2500-shape, C3 Ladybug-scale, C4 Venice-scale).
Checks: (1) the linear-solve delta of ONE damped trial against the unmodified reference's delta at full size
(tests/golden/fullsize__<w>_metis.npz, written by tests/golden/make_fullsize_golden.py: delta, both linear errors, the error
after the step; C4 keeps every 16th scalar) with the north_star bar 1e-9 relative -- or, where two FP64 factorisations
cannot agree that closely, the arbiter: delta_star = the same system solved to long-double accuracy by the reference
tool (conftest.check_delta); (2) two runs are bit-identical (the
hand-over into large fronts reduces in a fixed order); (3) properties that need no oracle: the linearised error at zero
is the nonlinear error, the step decreases the linearised error, the solution does not depend on the ordering."""
import json
import os

import numpy as np
import pytest

import bench
from conftest import GOLDEN, check_delta
from gtsam_petercdev_b200.engine import Engine

pytestmark = pytest.mark.gpu
REF = json.load(open(os.path.join(GOLDEN, "fullsize_trial_errors.json")))
LAM = 1e-5


def _trial(problem, workload, ordering):
    e = Engine(problem, device=0)
    e.set_ordering(bench.ordering_for(problem, workload, ordering))
    err0 = e.error()
    e.linearize()
    e.factorize(LAM)
    e.backsolve()
    lin0, lin1 = e.linear_errors()
    e.retract()
    return err0, lin0, lin1, e.error(1), e.delta()


def _rel(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / np.linalg.norm(b))


@pytest.mark.parametrize("workload", ["c1", "c2", "c3", "c4"])
def test_delta_at_full_size_matches_the_reference(workload):
    import make_golden
    g = dict(np.load(os.path.join(GOLDEN, f"fullsize__{workload}_metis.npz")))
    p = bench.make_problem(workload)
    assert make_golden.problem_digest(p) == str(g["digest"]), "full-size generator drifted from its fixture"
    err0, lin0, lin1, err1, delta = _trial(p, workload, "metis")
    assert len(delta) == int(g["n_scalars"][0])
    if workload == "c4":   # the fixture keeps every 16th scalar of delta (24 MB otherwise)
        d, ref, star = delta[::16], g["delta_stride16"], g["delta_star_stride16"]
    else:
        d, ref, star = delta, g["delta"], g["delta_star"]
    check_delta(f"fullsize {workload} metis lambda=1e-5", d, ref, star, f"   trial error rel = {abs(err1 - g['error1'][0]) / g['error1'][0]:.2e}")
    assert abs(err0 - g["error0"][0]) <= 1e-12 * g["error0"][0]
    assert abs(lin0 - g["linear_error0"][0]) <= 1e-12 * g["linear_error0"][0]
    assert abs(lin1 - g["linear_error1"][0]) <= 1e-8 * g["linear_error0"][0]
    assert abs(err1 - g["error1"][0]) <= 1e-9 * g["error1"][0]
    # the same trial again: every reduction runs in a fixed order, so the result repeats bit for bit
    _, _, _, again, delta2 = _trial(p, workload, "metis")
    assert again == err1 and np.array_equal(delta2, delta)


@pytest.mark.parametrize("workload", ["c1", "c2", "c3"])
def test_one_lm_trial_at_full_size_matches_the_reference(workload):
    p = bench.make_problem(workload)
    err0, lin0, lin1, err1, delta = _trial(p, workload, "metis")
    assert abs(err1 - REF[workload]) <= 1e-9 * REF[workload]            # the reference's number, computed on its CPU path
    assert abs(lin0 - err0) <= 1e-12 * err0                             # GaussianFactorGraph::error(0) == graph.error(values)
    assert lin1 < lin0 and np.all(np.isfinite(delta))


@pytest.mark.parametrize("workload,other", [("c1", "colamd"), ("c2", "colamd"), ("c3", "schur")])
def test_solution_does_not_depend_on_the_ordering(workload, other):
    """A different elimination tree (other cliques, other kernels classes, other summation order) must give the same
    damped Gauss-Newton step: delta to 1e-6 of its norm (these systems are ill-conditioned: the reference's own delta moves
    by up to 7e-7 under a one-ulp change of its inputs), the retracted error to 1e-9."""
    p = bench.make_problem(workload)
    _, _, lin1_a, err_a, d_a = _trial(p, workload, "metis")
    _, _, lin1_b, err_b, d_b = _trial(p, workload, other)
    assert np.linalg.norm(d_a - d_b) <= 1e-6 * np.linalg.norm(d_a)
    assert abs(lin1_a - lin1_b) <= 1e-9 * abs(lin1_a)
    assert abs(err_a - err_b) <= 1e-9 * err_a
    assert abs(err_b - REF[workload]) <= 1e-9 * REF[workload]


def test_c2_on_the_shipped_sphere2500():
    """C2 on the dataset the reference ships: examples/Data/sphere2500.txt through the native reader, the
    reference's own METIS ordering for that graph, one damped trial against the unmodified reference
    (tests/golden/fullsize__c2_real_metis.npz by tests/golden/make_c2_real_golden.py).  Same bars as the synthetic twin."""
    import make_golden
    from gtsam_petercdev_b200 import datasets
    g = dict(np.load(os.path.join(GOLDEN, "fullsize__c2_real_metis.npz")))
    p = datasets.sphere2500_file()
    assert make_golden.problem_digest(p) == str(g["digest"]), "the native reader / dead reckoning drifted from the fixture"
    keys = p.variables()[0]

    def trial(order):
        e = Engine(p, device=0)
        e.set_ordering(keys[g["ordering_" + order]])
        err0 = e.error()
        e.linearize(); e.factorize(LAM); e.backsolve()
        lin0, lin1 = e.linear_errors()
        e.retract()
        return err0, lin0, lin1, e.error(1), e.delta()
    err0, lin0, lin1, err1, delta = trial("metis")
    check_delta("fullsize c2_real (sphere2500.txt) metis", delta, g["delta"], g["delta_star"],
                f"   trial error rel = {abs(err1 - g['error1'][0]) / g['error1'][0]:.2e}")
    assert abs(err0 - g["error0"][0]) <= 1e-12 * g["error0"][0]
    assert abs(lin0 - g["linear_error0"][0]) <= 1e-12 * g["linear_error0"][0]
    assert abs(lin1 - g["linear_error1"][0]) <= 1e-8 * g["linear_error0"][0]
    assert abs(err1 - g["error1"][0]) <= 1e-9 * g["error1"][0]
    # the other ordering of the reference gives the same step
    _, _, _, err1c, deltac = trial("colamd")
    assert _rel(deltac, delta) < 1e-8 and abs(err1c - err1) <= 1e-9 * err1
