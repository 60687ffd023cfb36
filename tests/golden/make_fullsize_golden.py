"""Full-size fixtures (BASELINE.json configs C1, C2, C3, C4): ONE damped linear solve of the unmodified reference
(oracle/_ref/ref_tool mode=solve) at the generator's initial values, on the reference's own METIS ordering, lambda = 1e-5,
identity damping -- the trial bench.py times.  Stored: delta and delta_star = the same system solved to long-double accuracy (C1..C3 whole; C4 every 16th scalar + norms), both linear
errors, the error before and after the step.

    python tests/golden/make_fullsize_golden.py [c1 c2 c3 c4]

Needs /root/reference (this container only).  tests/test_fullsize_gpu.py compares the device against these.
"""
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench  # noqa: E402
import make_golden  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
LAM = 1e-5


def main():
    for w in (sys.argv[1:] or ["c1", "c2", "c3", "c4"]):
        p = bench.make_problem(w)
        keys = bench.ordering_for(p, w, "metis")
        p.a["ordering"] = np.ascontiguousarray(keys, dtype=np.uint64)
        f = tempfile.mktemp(suffix=".gtb")
        p.save(f)
        del p.a["ordering"]
        t = time.time()
        o = make_golden.run_tool(f, ["mode=solve", "ordering=file", f"lambda={LAM!r}", "diag=0", "tree=0", "ld=2"])
        os.remove(f)
        assert int(o["failed"][0]) == 0
        d = o["delta"]
        keep = dict(digest=make_golden.problem_digest(p), lam=LAM, error0=o["error0"], error1=o["error1"],
                    linear_error0=o["linear_error0"], linear_error1=o["linear_error1"], delta_norm=np.array([np.linalg.norm(d)]),
                    n_scalars=np.array([len(d)]), delta_star_sweeps=o["delta_star_sweeps"])
        # delta_star: the same damped system solved to long-double accuracy (ref_tool ld=2), the arbiter of the device's delta
        ds = o["delta_star"]
        keep["delta_ref_vs_star"] = np.array([np.linalg.norm(d - ds) / np.linalg.norm(ds)])
        keep["delta_star_norm"] = np.array([np.linalg.norm(ds)])
        if w == "c4":
            keep["delta_stride16"] = d[::16].copy()
            keep["delta_star_stride16"] = ds[::16].copy()
        else:
            keep["delta"] = d
            keep["delta_star"] = ds
        np.savez_compressed(os.path.join(OUT, f"fullsize__{w}_metis.npz"), **keep)
        print(w, "scalars", len(d), "error", float(o["error0"][0]), "->", float(o["error1"][0]), "|delta|", float(keep["delta_norm"][0]), "|d_ref - d*|/|d*|", float(keep["delta_ref_vs_star"][0]), "sweeps", o["delta_star_sweeps"],
              f"{time.time() - t:.1f} s")


if __name__ == "__main__":
    main()
