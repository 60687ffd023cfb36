"""Golden vectors of the ISAM2 row, made HERE by the unmodified reference (oracle/_ref/ref_tool
mode=isam2: gtsam::ISAM2 with the parameters and the update schedule of examples/ISAM2_City10000.cpp:62-134).

  tests/golden/data/T1_city10000_04_head1264.txt   the first 1264 lines (200 loop closures) of
                                                   examples/Data/T1_city10000_04.txt, so the input travels
  tests/golden/isam2__city10000_200loops.npz       estimate after every 100th update and at the end, counts, error

With --full it also runs the whole file (10,000 poses, 10,688 loop closures; 30 minutes on one core here) and stores how
far the result is from the ground-truth trajectory the reference ships for this experiment
(examples/Data/ISAM2_GT_city10000.txt: the noise-free city grid, not an ISAM2 output -- so a sanity bound, not a parity
pin), a stride-50 sample of the estimate, and the seconds of every update: the CPU baseline of the row.
usage: python tests/golden/make_isam2_golden.py [--full | --full-from <dump.gtb>]"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from gtsam_petercdev_b200 import gtbio  # noqa: E402

TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_tool")
HEAD = os.path.join(HERE, "data", "T1_city10000_04_head1264.txt")
FULL = "/root/reference/examples/Data/T1_city10000_04.txt"
GT = "/root/reference/examples/Data/ISAM2_GT_city10000.txt"


def run(data, **kw):
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "o.gtb")
        subprocess.run([TOOL, data, out, "mode=isam2"] + [f"{k}={v}" for k, v in kw.items()], check=True)
        return gtbio.load(out)


def main():
    d = run(HEAD, every=100)
    keep = {k: d[k] for k in ("final", "snapshots", "snapshot_update", "snapshot_nposes", "snapshot_offset", "counts",
                              "final_error", "update_is_loop")}
    keep["mean_update_seconds_here"] = np.array([d["update_seconds"].mean()])
    path = os.path.join(HERE, "isam2__city10000_200loops.npz")
    if "--full" in sys.argv or "--full-from" in sys.argv:
        # --full-from <out.gtb>: reuse the dump of an earlier whole-file run (30 minutes on one core here)
        f = gtbio.load(sys.argv[sys.argv.index("--full-from") + 1]) if "--full-from" in sys.argv else run(FULL, every=1000)
        gt = np.loadtxt(GT)
        est = f["final"]
        dist = np.linalg.norm(est[:, :2] - gt[:, :2], axis=1)
        keep["full_counts"] = f["counts"]
        keep["full_final_error"] = f["final_error"]
        keep["full_xy_distance_to_ground_truth_mean_max"] = np.array([dist.mean(), dist.max()])
        keep["full_update_seconds"] = f["update_seconds"].astype(np.float32)
        keep["full_final_stride50"] = est[::50].copy()
    elif os.path.exists(path):
        old = dict(np.load(path))
        for k in old:
            if k.startswith("full_"):
                keep[k] = old[k]
    np.savez_compressed(path, **keep)
    print("wrote", path, {k: v.shape for k, v in keep.items()})


if __name__ == "__main__":
    main()
