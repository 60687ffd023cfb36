"""The oracle of the ISAM2 row (the device side is tests/test_isam2_gpu.py): the reference's own
gtsam::ISAM2 driven by oracle/_ref/ref_tool mode=isam2 with the schedule of examples/ISAM2_City10000.cpp:62-134.
Checked here: the committed fixture against a fresh run of the reference (when it is built in this container),
its bookkeeping against an independent reading of the input file, the reported error against a numpy restatement, and
-- when the generator was run with --full -- the whole-file result against the ground-truth trajectory the reference
ships (examples/Data/ISAM2_GT_city10000.txt)."""
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(HERE, "golden", "isam2__city10000_200loops.npz")
HEAD = os.path.join(HERE, "golden", "data", "T1_city10000_04_head1264.txt")
TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_tool")


def read_city(path):
    rows = []
    for line in open(path):
        f = line.split()
        if len(f) < 9 or f[0] != "EDGE2":
            continue
        rows.append((int(f[1]), int(f[3]), int(f[5]), float(f[6]), float(f[7]), float(f[8])))
    return rows


def test_fixture_bookkeeping_matches_the_input_file():
    g = dict(np.load(GOLD))
    rows = read_city(HEAD)
    is_loop = np.array([0 if b == a + 1 else 1 for a, b, *_ in rows])
    updates, loops, nposes, nfactors = g["counts"].tolist()
    assert updates == len(rows) == 1264 and loops == int(is_loop.sum()) == 200
    assert np.array_equal(g["update_is_loop"], is_loop)
    assert nposes == max(b for a, b, *_ in rows if b == a + 1) + 1 == g["final"].shape[0]
    assert nfactors == updates + 1                                    # one factor per line + the prior
    # snapshots: every 100th update, all poses known at that time
    assert g["snapshot_update"].tolist() == list(range(100, updates + 1, 100))
    known = np.cumsum(1 - is_loop) + 1
    assert g["snapshot_nposes"].tolist() == [int(known[u - 1]) for u in g["snapshot_update"]]
    assert g["snapshot_offset"].tolist() == [0] + np.cumsum(3 * g["snapshot_nposes"])[:-1].tolist()
    assert g["snapshots"].size == 3 * int(g["snapshot_nposes"].sum())
    # pose 0 stays on its 1e-4 prior; the first step is the first measurement
    assert np.abs(g["final"][0]).max() < 1e-6
    assert np.allclose(g["final"][1], rows[0][3:], atol=5e-3)
    assert np.isfinite(g["final"]).all() and float(g["final_error"][0]) > 0.0


def between(a, b):
    """a^-1 b for Pose2 rows (x, y, theta) (geometry/Pose2.h:138-140, Pose2.cpp:202-204)."""
    c, s = np.cos(a[..., 2]), np.sin(a[..., 2])
    dx, dy = b[..., 0] - a[..., 0], b[..., 1] - a[..., 1]
    th = np.angle(np.exp(1j * (b[..., 2] - a[..., 2])))
    return np.stack([c * dx + s * dy, -s * dx + c * dy, th], axis=-1)


def test_error_of_the_estimate_restated_in_numpy():
    """Independent restatement of the graph the oracle built: 0.5 * sum |Sigma^-1/2 Local(z, x_i^-1 x_j)|^2 over the
    BetweenFactor<Pose2> of every line (slam/BetweenFactor.h:111-124, first-order Pose2 chart, geometry/Pose2.cpp:112-122)
    with the example's noise models, plus the prior: must equal the error the reference reports for its own estimate."""
    g = dict(np.load(GOLD))
    x = g["final"]
    rows = np.array(read_city(HEAD))
    a, b = rows[:, 0].astype(int), rows[:, 1].astype(int)
    z = rows[:, 3:6]
    e = between(z, between(x[a], x[b]))
    odo = b == a + 1
    sig = np.where(odo[:, None], np.array([1 / 30.0, 1 / 30.0, 1 / 100.0])[None, :], 10.0)
    total = 0.5 * np.sum((e / sig) ** 2) + 0.5 * np.sum((x[0] / 1e-4) ** 2)
    assert abs(total - float(g["final_error"][0])) <= 1e-9 * float(g["final_error"][0])
    # and the estimate is far better than dead reckoning, which ignores the loop closures
    dr = np.zeros_like(x)
    for ai, bi, zi in zip(a, b, z):
        if bi == ai + 1:
            c, s = np.cos(dr[ai, 2]), np.sin(dr[ai, 2])
            dr[bi] = (dr[ai, 0] + c * zi[0] - s * zi[1], dr[ai, 1] + s * zi[0] + c * zi[1], dr[ai, 2] + zi[2])
    e_dr = between(z, between(dr[a], dr[b]))
    assert 0.5 * np.sum((e_dr / sig) ** 2) > 2.0 * total


@pytest.mark.skipif(not os.path.exists(TOOL), reason="oracle/_ref is built only where /root/reference exists")
def test_reference_reproduces_the_fixture(tmp_path):
    import sys
    sys.path.insert(0, ROOT)
    from gtsam_petercdev_b200 import gtbio
    out = str(tmp_path / "o.gtb")
    subprocess.run([TOOL, HEAD, out, "mode=isam2", "every=100"], check=True, timeout=300)
    d = gtbio.load(out)
    g = dict(np.load(GOLD))
    for k in ("counts", "snapshot_update", "snapshot_nposes", "snapshot_offset", "update_is_loop"):
        assert np.array_equal(d[k], g[k]), k
    for k in ("final", "snapshots", "final_error"):
        assert np.allclose(d[k], g[k], rtol=0, atol=1e-12), k       # serial reference: deterministic


def test_whole_file_run_stays_on_the_ground_truth():
    """examples/Data/ISAM2_GT_city10000.txt is the ground-truth trajectory the reference ships for this experiment (the
    noise-free grid, not an ISAM2 output): after all 20,687 updates the reference's estimate is within a cell of it
    everywhere, where dead reckoning over the same odometry ends up 33 m off on average."""
    g = dict(np.load(GOLD))
    if "full_counts" not in g:
        pytest.skip("generator was not run with --full")
    updates, loops, nposes, nfactors = g["full_counts"].tolist()
    assert nposes == 10000 and updates == 20687 and loops == 10688 and nfactors == updates + 1
    mean, worst = g["full_xy_distance_to_ground_truth_mean_max"].tolist()
    assert mean < 1.0 and worst < 3.0
    us = g["full_update_seconds"]
    assert us.size == updates and (us > 0).all()
