"""Several engines, one per rank, against the single-engine result and the reference.

The ranks here are threads of one process sharing cuda:0 and the collectives are gtsam_petercdev_b200.distributed.ThreadComm;
the library sees exactly the callbacks it gets from NCCL under torchrun (bench.py --gpus N).  What is checked: the tree
cut, the fan-in of the top fronts, the broadcast / owned-delta sum of back-substitution, and that LM takes the same
decisions on every rank."""
import os
import threading

import numpy as np
import pytest

from conftest import golden_problem, load_golden, run_tags, solve_tags
from gtsam_petercdev_b200.distributed import ThreadComm
from gtsam_petercdev_b200.engine import Engine

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300)


def run_ranks(world, body):
    comm = ThreadComm(world)
    out, err = [None] * world, [None] * world

    def work(r):
        try:
            out[r] = body(r, comm.callback(r))
        except Exception as ex:   # noqa: BLE001
            err[r] = ex
            comm.barrier.abort()

    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join(300) for t in th]
    for e in err:
        if e is not None:
            raise e
    return out, comm


@pytest.fixture
def force_top(monkeypatch):
    monkeypatch.setenv("GTB_MIN_TOP", "2")


SOLVES = [t for t in solve_tags() if t[1] in ("metis", "colamd")]


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("name,ordering,lam,diag,tag", SOLVES)
def test_fan_in_solve_matches_reference(world, name, ordering, lam, diag, tag, force_top):
    g = load_golden(tag)
    p = golden_problem(name)

    def body(r, cb):
        e = Engine(p, device=0)
        e.set_distributed(r, world, cb)
        e.set_ordering(g["ordering"])
        part = e.partition()
        e.linearize()
        e.factorize(lam, bool(diag))
        e.backsolve()
        d = e.delta()
        e0, e1 = e.linear_errors()
        e.retract()
        return dict(delta=d, e1=e1, err=e.error(1), part=part)

    out, comm = run_ranks(world, body)
    sens = float(g["delta_ulp_sensitivity"][0])
    for r in range(world):
        assert np.array_equal(out[r]["part"]["owner"], out[0]["part"]["owner"])
        assert np.array_equal(out[r]["delta"], out[0]["delta"])          # replicas hold the same bits
        assert rel(out[r]["delta"], g["delta"]) < max(1e-9, 3 * sens)
        assert abs(out[r]["err"] - g["error1"][0]) <= max(1e-9, 10 * sens) * max(g["error1"][0], g["error0"][0] * 1e-3)
    owner = out[0]["part"]["owner"]
    if out[0]["part"]["n_top"]:
        assert comm.calls[0] == 1 and comm.calls[1] == 1     # one fan-in, one broadcast
        assert set(np.unique(owner[owner >= 0])) <= set(range(world))
    assert comm.calls[2] == 1                                 # owned pieces of delta summed once


@pytest.mark.parametrize("name,mode,ordering,extra,tag", [t for t in run_tags() if t[1] == "lm" and t[2] == "metis"])
def test_lm_takes_the_same_path_on_every_rank(name, mode, ordering, extra, tag, force_top):
    g = load_golden(tag)
    p = golden_problem(name)

    def body(r, cb):
        e = Engine(p, device=0)
        e.set_distributed(r, 2, cb)
        e.set_ordering(g["ordering"])
        return e.lm_optimize(e.lm_params(ceres="ceres=1" in extra))

    out, _ = run_ranks(2, body)
    for r in range(2):
        assert out[r]["iterations"] == int(g["iterations"][0])
        assert out[r]["inner_iterations"] == int(g["inner_iterations"][0])
        assert abs(out[r]["final_error"] - g["final_error"][0]) <= 1e-6 * g["final_error"][0]
    assert out[0]["final_error"] == out[1]["final_error"]


def test_a_failed_factorisation_is_agreed_across_ranks(force_top):
    """A whole trial (gtb_trial) carries the failure flag on the delta all-reduce; when any rank's factorisation fails, every
    rank raises IndeterminantLinearSystemException with the same key, and a damped trial afterwards succeeds everywhere."""
    from gtsam_petercdev_b200.engine import IndeterminantLinearSystemException
    from gtsam_petercdev_b200.problem import Problem
    p0 = golden_problem("ba_60_2000")
    a = {k: v for k, v in p0.a.items() if not k.startswith("prior")}   # gauge freedom: undamped elimination must fail
    g = load_golden("ba_60_2000__solve_metis_0.0001_1")

    def body(r, cb):
        e = Engine(Problem(a), device=0)
        e.set_distributed(r, 2, cb)
        e.set_ordering(g["ordering"])
        e.linearize()
        key = None
        try:
            e.trial(0.0)
        except IndeterminantLinearSystemException as ex:
            key = ex.key
        ok = e.trial(1e-3)
        return key, ok

    out, _ = run_ranks(2, body)
    assert out[0][0] is not None and out[0][0] == out[1][0]
    assert out[0][1] == out[1][1] and np.all(np.isfinite(out[0][1]))
