"""Host logic of the static launch schedule (the replacement of the reference's task-parallel tree walk,
inference/ClusterTree-inst.h:286-318) on CPU: gtb_schedule_check verifies inside the library that every clique is
eliminated and solved exactly once, that every member of a launch class fits the class's kernel and shared memory,
and that the task lists, inverse blocks and camera-pair segments tile their buffers.  Integer work: exact."""
import numpy as np
import pytest

from conftest import golden_problem, load_golden, solve_tags
from gtsam_petercdev_b200.engine import Engine


@pytest.mark.parametrize("name,ordering,lam,diag,tag", solve_tags())
def test_schedule_invariants(name, ordering, lam, diag, tag):
    e = Engine(golden_problem(name), device=-1)
    e.set_ordering(load_golden(tag)["ordering"])
    st = e.schedule_check()
    pl = e.plan()
    assert st["launch_classes"] >= 1 and st["backsolve_launches"] >= int(pl["stats"]["levels"])
    if name.startswith("ba_"):
        # every projection factor of a point clique appears once in a same-camera list
        assert st["diag_segments"] >= 1
    else:
        assert st["pair_segments"] == 0 and st["diag_segments"] == 0
    # large fronts exist exactly when there are panel steps, Schur steps and inverse blocks to schedule
    n_large = int(pl["stats"]["n_large"])
    assert (st["panel_step_launches"] > 0) == (n_large > 0)


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_schedule_invariants_of_every_rank(world, monkeypatch):
    monkeypatch.setenv("GTB_MIN_TOP", "2")
    g = load_golden("ba_60_2000__solve_metis_0.0001_1")
    steps = 0
    for rank in range(world):
        e = Engine(golden_problem("ba_60_2000"), device=-1)
        e.set_distributed(rank, world)
        e.set_ordering(g["ordering"])
        st = e.schedule_check()           # both phases: own subtrees, and the shared top
        steps += st["step_tasks"]
    assert steps > 0


@pytest.mark.parametrize("env,value", [("GTB_POINT_ATOMIC", "1"), ("GTB_PAIR_SEG", "5"), ("GTB_DIAG_SEG", "3")])
def test_schedule_invariants_under_switches(env, value, monkeypatch):
    """The experiment switches change list shapes, never coverage."""
    order = load_golden("ba_30_500__solve_metis_0.0001_1")["ordering"]
    e0 = Engine(golden_problem("ba_30_500"), device=-1)
    e0.set_ordering(order)
    base = e0.schedule_check()
    monkeypatch.setenv(env, value)
    e = Engine(golden_problem("ba_30_500"), device=-1)
    e.set_ordering(order)
    st = e.schedule_check()
    if env == "GTB_POINT_ATOMIC":
        assert st["pair_segments"] == 0 and st["diag_segments"] == 0 and base["diag_segments"] > 0
    elif env == "GTB_PAIR_SEG":
        assert st["pair_segments"] > base["pair_segments"] and st["diag_segments"] == base["diag_segments"]
    else:
        assert st["diag_segments"] > base["diag_segments"] and st["pair_segments"] == base["pair_segments"]


def test_schedule_check_needs_a_plan():
    from gtsam_petercdev_b200.engine import EngineError
    e = Engine(golden_problem("pose2_300"), device=-1)
    with pytest.raises(EngineError):
        e.schedule_check()


@pytest.mark.parametrize("world", [2, 8])
def test_schedule_invariants_at_full_size(world):
    """C3 (Ladybug-scale, the reference's METIS ordering): every rank's schedule of a `world`-GPU cut, and the ranks'
    subtree cliques together with the shared top cover the tree exactly once."""
    import bench
    p = bench.make_problem("c3")
    keys = bench.ordering_for(p, "c3", "metis")
    covered = None
    for rank in range(world):
        e = Engine(p, device=-1)
        e.set_distributed(rank, world)
        e.set_ordering(keys)
        st = e.schedule_check()
        assert st["panel_step_launches"] > 0 and st["diag_segments"] > 0
        owner = e.partition()["owner"]
        mine = (owner == rank).astype(np.int64)
        covered = mine if covered is None else covered + mine
    covered += (owner == -1)
    assert covered.min() == 1 and covered.max() == 1
