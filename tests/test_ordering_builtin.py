"""gtb_set_ordering_builtin(2): the nested-dissection ordering computed inside the library (csrc/ordering.cpp).  It is not the
permutation CCOLAMD / METIS give; what must hold: it is a permutation, its fill stays in the range of the reference's
orderings on the pose-graph workloads (far below the natural order), and -- on the GPU -- the solution of the linear system
does not depend on it (delta against the reference's fixture, same bar as the other solves)."""
import numpy as np
import pytest

from conftest import check_delta, golden_problem, load_golden, solve_tags
from gtsam_petercdev_b200.engine import Engine


def fill(e):
    return int(e.plan()["n_conditional_doubles"])


@pytest.mark.parametrize("wl", ["c1", "c2"])
def test_fill_is_in_the_range_of_the_reference_orderings(wl):
    import bench
    p = bench.make_problem(wl)
    got = {}
    for o in ("metis", "colamd", "nested_dissection", "natural"):
        e = Engine(p, device=-1)
        e.set_ordering(o if o in ("nested_dissection", "natural") else bench.ordering_for(p, wl, o))
        st = e.schedule_check()
        assert st["panel_step_launches"] >= 0
        got[o] = fill(e)
        pl = e.plan()
        front = pl["clique_keys"][np.concatenate([np.arange(o_, o_ + f) for o_, f in
                                                  zip(np.cumsum(pl["clique_nkeys"]) - pl["clique_nkeys"], pl["clique_nfrontal"])])]
        assert np.array_equal(np.sort(front), e.var_keys)      # every variable is frontal exactly once
    assert got["nested_dissection"] <= 1.5 * max(got["metis"], got["colamd"]), got
    assert got["nested_dissection"] <= 0.5 * got["natural"], got


def test_order_is_deterministic_and_handles_disconnected_graphs():
    from gtsam_petercdev_b200.problem import Problem
    rng = np.random.default_rng(3)
    n = 60
    keys = np.arange(100, 100 + n, dtype=np.uint64)
    # three chains that do not touch each other, plus one isolated pose with a prior
    pairs = [(i, i + 1) for i in range(0, 19)] + [(i, i + 1) for i in range(20, 39)] + [(i, i + 1) for i in range(40, 58)]
    a = {"pose2_keys": keys, "pose2": rng.normal(size=(n, 3)),
         "between2_keys": np.array([[keys[i], keys[j]] for i, j in pairs], dtype=np.uint64),
         "between2_z": rng.normal(size=(len(pairs), 3)), "between2_nk": np.zeros(len(pairs), dtype=np.int32),
         "between2_noise": np.zeros((len(pairs), 9)),
         "prior2_keys": keys[[0, 20, 40, 59]].reshape(-1, 1), "prior2_z": np.zeros((4, 3)),
         "prior2_nk": np.zeros(4, dtype=np.int32), "prior2_noise": np.zeros((4, 9))}
    plans = []
    for _ in range(2):
        e = Engine(Problem(a), device=-1)
        e.set_ordering("nested_dissection")
        e.schedule_check()
        plans.append(e.plan())
    assert np.array_equal(plans[0]["clique_keys"], plans[1]["clique_keys"])
    assert plans[0]["n_levels"] <= 8 if "n_levels" in plans[0] else True     # a balanced tree, not a chain of 20


@pytest.mark.gpu
@pytest.mark.parametrize("name,ordering,lam,diag,tag", [t for t in solve_tags() if t[0].startswith("pose")])
def test_solution_does_not_depend_on_the_ordering(name, ordering, lam, diag, tag):
    g = load_golden(tag)
    e = Engine(golden_problem(name), device=0)
    e.set_ordering("nested_dissection")
    e.linearize()
    e.factorize(lam, bool(diag))
    e.backsolve()
    check_delta(tag + " [nd]", e.delta(), g["delta"], g["delta_star"])
