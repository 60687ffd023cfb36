"""ISAM2 row on the device: gtb_isam2_update driven with the schedule of
examples/ISAM2_City10000.cpp:62-134 (prior on pose 0, then one update per measurement line, relinearizeThreshold 0.01,
relinearizeSkip 1, Gauss-Newton with wildfire threshold 0) against the estimates the reference's own gtsam::ISAM2 produced
for the same 1264 lines (tests/golden/isam2__city10000_200loops.npz, made by tests/golden/make_isam2_golden.py).
Tolerance: 1e-9 absolute on x, y and theta of every pose at every snapshot (north_star's 1e-9)."""
import os

import numpy as np
import pytest

from gtsam_petercdev_b200.isam2 import DIAG_SIGMAS, ISAM2, X, drive_city10000 as drive, read_city10000 as read_city

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden", "isam2__city10000_200loops.npz")
HEAD = os.path.join(HERE, "golden", "data", "T1_city10000_04_head1264.txt")
TOL = 1e-9


def xyt(rows):
    """device rows (cos, sin, x, y) -> (x, y, theta)"""
    return np.stack([rows[:, 2], rows[:, 3], np.arctan2(rows[:, 1], rows[:, 0])], axis=1)


def angle_diff(a, b):
    return np.angle(np.exp(1j * (a - b)))


def test_estimates_after_every_update_match_the_reference():
    g = dict(np.load(GOLD))
    rows = read_city(HEAD)
    snaps = {int(u): (int(n), int(o)) for u, n, o in zip(g["snapshot_update"], g["snapshot_nposes"], g["snapshot_offset"])}
    worst = {}

    def check(u, est, row_of):
        if u not in snaps:
            return
        n, off = snaps[u]
        want = g["snapshots"][off:off + 3 * n].reshape(n, 3)
        got = xyt(est)[[row_of[i] for i in range(n)]]
        d = np.abs(got - want)
        d[:, 2] = np.abs(angle_diff(got[:, 2], want[:, 2]))
        worst[u] = float(d.max())

    isam = ISAM2(relinearize_threshold=0.01, relinearize_skip=1)
    est, row_of, results = drive(isam, rows, check)
    n = g["final"].shape[0]
    got = xyt(est)[[row_of[i] for i in range(n)]]
    d = np.abs(got - g["final"])
    d[:, 2] = np.abs(angle_diff(got[:, 2], g["final"][:, 2]))
    worst["final"] = float(d.max())
    print("\nISAM2 parity, max |estimate - reference| per snapshot:", {k: "%.2e" % v for k, v in worst.items()})
    from conftest import record_parity
    record_parity(f"{'isam2 city10000 head, 1264 updates':36s} max |estimate - reference| over 12 snapshots + final = "
                  f"{max(worst.values()):.3e}   (bar {TOL:g} absolute on x, y, theta)")
    assert sorted(k for k in worst if k != "final") == sorted(snaps)
    assert max(worst.values()) < TOL, worst
    # bookkeeping of the reference run: updates, poses, factors (one per line + the prior)
    updates, loops, nposes, nfactors = g["counts"].tolist()
    last = results[-1]
    assert last["update_count"] == updates + 1 and last["n_variables"] == nposes and last["n_factors"] == nfactors
    # fluid relinearisation did something, and not everything every time
    moved = np.array([r["variables_relinearized"] for r in results])
    assert moved.max() > 0 and (moved == 0).sum() > 0
    isam.close()


def test_estimate_does_not_depend_on_the_elimination_order():
    """The reference orders every update with constrained COLAMD (ISAM2.cpp:301-340); any order must give its estimate."""
    rows = read_city(HEAD)[:260]
    a = ISAM2(0.01, 1)
    est_a, row_a, _ = drive(a, rows)
    b = ISAM2(0.01, 1)
    rng = np.random.default_rng(5)
    n = max(r[1] for r in rows if r[1] == r[0] + 1) + 1
    b.set_ordering([X(i) for i in rng.permutation(n)])
    est_b, row_b, _ = drive(b, rows)
    assert row_a == row_b
    d = np.abs(xyt(est_a) - xyt(est_b))
    d[:, 2] = np.abs(angle_diff(xyt(est_a)[:, 2], xyt(est_b)[:, 2]))
    assert d.max() < 1e-9
    a.close(); b.close()


def test_relinearize_marked_moves_exactly_the_variables_over_the_threshold():
    """gtb_relinearize_marked (ISAM2.cpp:456-468, ISAM2-impl.h:341-351) on a batch engine: rows with |delta|_inf >= beta take
    the retracted value bit for bit, the others keep the current one."""
    import ctypes as C
    from conftest import golden_problem
    from gtsam_petercdev_b200.engine import Engine, _ptr
    e = Engine(golden_problem("pose2_300"), device=0)
    e.set_ordering("natural")
    e.linearize(); e.factorize(0.0); e.backsolve(); e.retract()
    cur, new = e.variables("pose2", 0), e.variables("pose2", 1)
    delta = e.delta().reshape(-1, 3)           # ascending keys = upload order for this problem
    assert np.array_equal(np.argsort(e.var_keys), np.arange(len(e.var_keys)))
    beta = float(np.median(np.abs(delta).max(axis=1)))
    want = np.abs(delta).max(axis=1) >= beta
    marks = np.zeros(len(cur), dtype=np.uint8)
    n = C.c_int64(0)
    e._ck(e.lib.gtb_relinearize_marked(e.h, beta, _ptr(marks), C.byref(n)))
    assert n.value == int(want.sum()) and 0 < n.value < len(cur)
    assert np.array_equal(marks.astype(bool), want)
    after = e.variables("pose2", 0)
    assert np.array_equal(after[want], new[want]) and np.array_equal(after[~want], cur[~want])
    with pytest.raises(Exception):
        e.backsolve()                         # the factorisation belongs to the old linearisation point


def test_underdetermined_update_raises_and_the_next_one_recovers():
    """ISAM2::update throws IndeterminantLinearSystemException from the elimination when the gauge is free (after the factors
    were pushed: they stay in the graph); a later update that fixes the gauge succeeds on the same object."""
    from gtsam_petercdev_b200.engine import IndeterminantLinearSystemException
    isam = ISAM2(0.01, 1)
    isam.add_variables("pose2", [X(0), X(1)], [[0.0, 0.0, 0.0], [1.0, 0.0, 0.0]])
    isam.add_factors("between2", [[X(0), X(1)]], [[1.0, 0.0, 0.0]], DIAG_SIGMAS, [[0.1, 0.1, 0.05]])
    with pytest.raises(IndeterminantLinearSystemException) as ei:
        isam.update()
    assert ei.value.key in (X(0), X(1))
    assert isam.count("pose2") == 2
    with pytest.raises(Exception):
        isam.calculate_estimate("pose2")          # no solution yet
    isam.add_factors("prior2", [X(0)], [[0.0, 0.0, 0.0]], DIAG_SIGMAS, [[1e-3, 1e-3, 1e-3]])
    r = isam.update()
    assert r["n_factors"] == 2 and r["n_variables"] == 2 and r["variables_relinearized"] == 0
    est = xyt(isam.calculate_estimate("pose2"))
    assert np.abs(est - np.array([[0.0, 0.0, 0.0], [1.0, 0.0, 0.0]])).max() < 1e-9
    isam.close()
