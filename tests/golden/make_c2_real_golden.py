"""C2 on the dataset the reference ships ("use the shipped examples/Data/sphere2500.txt"): the graph of
examples/Pose3SLAMExample_g2o.cpp:42-55 read by the NATIVE reader (datasets.sphere2500_file), the reference's own METIS and
COLAMD orderings for it, and one damped trial of the unmodified reference at the file's initial values (ref_tool mode=solve,
lambda = 1e-5, identity damping) with delta_star = the same system solved to long-double accuracy.

    python tests/golden/make_c2_real_golden.py      ->  tests/golden/fullsize__c2_real_metis.npz

Needs oracle/_ref (this container).  tests/test_fullsize_gpu.py::test_c2_on_the_shipped_sphere2500 compares the device."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden  # noqa: E402
from gtsam_petercdev_b200 import datasets  # noqa: E402

LAM = 1e-5


def main():
    p = datasets.sphere2500_file()
    keys = p.variables()[0]
    f = tempfile.mktemp(suffix=".gtb")
    p.save(f)
    keep = dict(digest=make_golden.problem_digest(p), lam=LAM)
    for o in ("metis", "colamd"):
        r = make_golden.run_tool(f, ["mode=ordering", f"ordering={o}"])
        perm = np.searchsorted(keys, r["ordering"]).astype(np.int32)
        assert np.array_equal(keys[perm], r["ordering"]) and len(np.unique(perm)) == len(keys)
        keep["ordering_" + o] = perm
    p.a["ordering"] = np.ascontiguousarray(keys[keep["ordering_metis"]], dtype=np.uint64)
    p.save(f)
    o = make_golden.run_tool(f, ["mode=solve", "ordering=file", f"lambda={LAM!r}", "diag=0", "tree=0", "ld=2"])
    os.remove(f)
    assert int(o["failed"][0]) == 0
    d, ds = o["delta"], o["delta_star"]
    keep.update(error0=o["error0"], error1=o["error1"], linear_error0=o["linear_error0"], linear_error1=o["linear_error1"],
                delta=d, delta_star=ds, n_scalars=np.array([len(d)]),
                delta_ref_vs_star=np.array([np.linalg.norm(d - ds) / np.linalg.norm(ds)]))
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "fullsize__c2_real_metis.npz"), **keep)
    print("c2_real: poses", len(keys), "scalars", len(d), "error", float(o["error0"][0]), "->", float(o["error1"][0]),
          "|d_ref - d*|/|d*|", float(keep["delta_ref_vs_star"][0]))


if __name__ == "__main__":
    main()
