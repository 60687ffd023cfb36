"""Host side of the multi-GPU path on CPU: the tree cut every rank computes for itself, checked between two
real processes over gloo, and the exact-sum reconstruction of delta from owned pieces."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _partition(problem_name, ordering_name, rank, world, min_top=None):
    from conftest import golden_problem, load_golden
    from gtsam_petercdev_b200.engine import Engine
    if min_top is not None:
        os.environ["GTB_MIN_TOP"] = str(min_top)
    g = load_golden(ordering_name)
    e = Engine(golden_problem(problem_name), device=-1)
    e.set_distributed(rank, world)          # no device: no collectives needed, only the plan
    e.set_ordering(g["ordering"])
    return e, e.partition(), e.plan()


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_tree_cut_is_a_partition(world, monkeypatch):
    monkeypatch.setenv("GTB_MIN_TOP", "2")
    e, part, plan = _partition("ba_60_2000", "ba_60_2000__solve_metis_0.0001_1", 0, world)
    owner, parent = part["owner"], plan["clique_parent"]
    assert owner.min() >= -1 and owner.max() < world
    # n_top counts the fronts the device factorises at the top (reference cliques merged by relaxed amalgamation count once)
    assert 2 <= part["n_top"] <= int((owner == -1).sum())
    for c in range(len(owner)):
        if parent[c] >= 0:
            if owner[c] == -1:
                assert owner[parent[c]] == -1                 # the top is closed towards the roots
            if owner[parent[c]] >= 0:
                assert owner[c] == owner[parent[c]]           # subtrees are never split
    shares = part["rank_share"][:world]
    assert abs(part["top_share"] + shares.sum() - 1.0) < 1e-12
    # the rank argument does not change the map
    _, part1, _ = _partition("ba_60_2000", "ba_60_2000__solve_metis_0.0001_1", world - 1, world)
    assert np.array_equal(part1["owner"], owner)


def test_one_rank_owns_everything():
    _, part, _ = _partition("pose2_300", "pose2_300__solve_metis_0.001_0", 0, 1)
    assert part["n_top"] == 0 and np.all(part["owner"] == 0)


def test_numeric_calls_need_collectives_or_a_device():
    from gtsam_petercdev_b200.engine import EngineError
    e, _, _ = _partition("pose2_300", "pose2_300__solve_metis_0.001_0", 1, 2)
    with pytest.raises(EngineError):
        e.linearize()                                          # no device: no CPU fallback
    with pytest.raises(EngineError):
        e.set_distributed(2, 2)                                # rank out of range


def _gloo_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), GTB_MIN_TOP="2")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        e, part, plan = _partition("ba_60_2000", "ba_60_2000__solve_metis_0.0001_1", rank, world)
        owner = torch.from_numpy(part["owner"].astype(np.int64))
        got = [torch.zeros_like(owner) for _ in range(world)]
        dist.all_gather(got, owner)
        same = all(torch.equal(t, owner) for t in got)
        # delta: each rank fills the scalars it owns (frontals of its cliques; the top belongs to rank 0), zeros elsewhere;
        # the sum over ranks must rebuild the vector bit for bit (own_delta_kernel + all-reduce on the device)
        rng = np.random.default_rng(7)
        truth = rng.standard_normal(plan["n_scalars"])
        keys, dims = e.var_keys, e.var_dim
        off = np.concatenate([[0], np.cumsum(dims)])
        mine = np.zeros(plan["n_scalars"], dtype=bool)
        kpos = np.concatenate([[0], np.cumsum(plan["clique_nkeys"])])
        for c, (nf, o) in enumerate(zip(plan["clique_nfrontal"], part["owner"])):
            if o == rank or (o == -1 and rank == 0):
                for k in plan["clique_keys"][kpos[c]:kpos[c] + nf]:
                    v = int(np.searchsorted(keys, k))
                    mine[off[v]:off[v + 1]] = True
        piece = torch.from_numpy(np.where(mine, truth, 0.0))
        count = torch.from_numpy(mine.astype(np.int64))
        dist.all_reduce(piece)
        dist.all_reduce(count)
        q.put((rank, same, bool(np.array_equal(piece.numpy(), truth)), bool((count == 1).all()), int(mine.sum())))
    finally:
        dist.destroy_process_group()


def test_two_processes_agree_over_gloo():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted(q.get(timeout=180) for _ in procs)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank, same, exact, once, n_mine in res:
        assert same, "ranks computed different tree cuts"
        assert exact, "sum of owned delta pieces is not the full vector"
        assert once, "a scalar has zero or several owners"
        assert n_mine > 0
