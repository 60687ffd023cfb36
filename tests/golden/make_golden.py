"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference (oracle/_ref/ref_tool,
built from /root/reference by oracle/Makefile) on seeded synthetic problems.

    make -C oracle -j8 ref && python tests/golden/make_golden.py

Needs /root/reference (this container only); the .npz files it writes are committed and travel to the GPU box.
Each fixture stores the oracle's outputs plus a checksum of the generated problem, so a test can regenerate the
problem from its seed and detect any drift.
"""
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from gtsam_petercdev_b200 import datasets, gtbio  # noqa: E402

TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_tool")
OUT = os.path.dirname(os.path.abspath(__file__))

# name -> (generator, kwargs)
PROBLEMS = {
    "pose2_300": (datasets.m3500_shape, dict(n_poses=300, n_closures=100, seed=3)),
    "pose3_120": (datasets.sphere_shape, dict(rings=10, per_ring=12, seed=5)),
    "ba_30_500": (datasets.ba_shape, dict(n_cams=30, n_points=500, mean_obs=4.0, window=10, seed=1)),
    "ba_60_2000": (datasets.ba_shape, dict(n_cams=60, n_points=2000, mean_obs=4.5, window=40, seed=7)),
}
# (problem, ordering, lambda, diag)
SOLVES = [
    ("pose2_300", "colamd", 0.0, 0), ("pose2_300", "metis", 1e-3, 0), ("pose2_300", "natural", 1e-2, 1),
    ("pose3_120", "colamd", 0.0, 0), ("pose3_120", "metis", 1e-3, 1), ("pose3_120", "natural", 1.0, 0),
    ("ba_30_500", "colamd", 1e-5, 0), ("ba_30_500", "metis", 1e-4, 1), ("ba_30_500", "schur", 1e-2, 0),
    ("ba_60_2000", "colamd", 1e-2, 0), ("ba_60_2000", "metis", 1e-4, 1), ("ba_60_2000", "schur", 1e-3, 1),
]
# (problem, mode, ordering, extra args)
RUNS = [
    ("pose2_300", "lm", "colamd", []), ("pose2_300", "gn", "colamd", []),
    ("pose3_120", "lm", "metis", []), ("pose3_120", "gn", "colamd", []),
    ("ba_30_500", "lm", "colamd", []), ("ba_30_500", "lm", "schur", ["ceres=1"]),
    ("ba_60_2000", "lm", "metis", ["ceres=1"]),
]


def problem_digest(p):
    h = hashlib.sha256()
    for k in sorted(p.a):
        h.update(k.encode())
        h.update(np.ascontiguousarray(p.a[k]).tobytes())
    return h.hexdigest()


def make_problem(name):
    gen, kw = PROBLEMS[name]
    return gen(**kw)


def run_tool(pfile, args):
    with tempfile.NamedTemporaryFile(suffix=".gtb", delete=False) as f:
        out = f.name
    try:
        subprocess.run([TOOL, pfile, out] + args, check=True)
        return gtbio.load(out)
    finally:
        if os.path.exists(out):
            os.remove(out)


def main():
    if not os.path.exists(TOOL):
        raise SystemExit("build the oracle first: make -C oracle -j8 ref")
    tmp = tempfile.mkdtemp()
    pfiles, digests = {}, {}
    for name in PROBLEMS:
        p = make_problem(name)
        pfiles[name] = os.path.join(tmp, name + ".gtb")
        p.save(pfiles[name])
        digests[name] = problem_digest(p)
        if "--solves-only" in sys.argv:
            continue
        lin = run_tool(pfiles[name], ["mode=linearize"])
        np.savez_compressed(os.path.join(OUT, f"{name}__linearize.npz"), digest=digests[name], **lin)
        print(name, "linearize", {k: v.shape for k, v in lin.items()})
    # conditioning yardstick: the reference's OWN delta moves by this much when every variable is perturbed by one
    # unit in the last place; no implementation with a different rounding sequence can agree more closely than that
    ulp_files = {}
    rng = np.random.default_rng(12345)
    for name in PROBLEMS:
        q = make_problem(name)
        for k in ("pose2", "pose3", "cam", "point"):
            if k in q.a:
                q.a[k] = np.nextafter(q.a[k], q.a[k] + rng.choice([-1.0, 1.0], size=q.a[k].shape))
        ulp_files[name] = os.path.join(tmp, name + "_ulp.gtb")
        q.save(ulp_files[name])
    for name, ordering, lam, diag in SOLVES:
        args = ["mode=solve", f"ordering={ordering}", f"lambda={lam!r}", f"diag={diag}"]
        # ld=1: also delta_star, the same damped system solved in 80-bit long double (the arbiter both the reference's
        # and the device's delta are measured against, oracle/ref_tool.cpp solve_long_double)
        # delta_star: by long-double iterative refinement on the reference's own factorisation (ld2=1: converges to ~1e-16, any
        # size); delta_star_dense: dense long-double normal equations (ld=1: independent check, resolution cond * 5e-20)
        o = run_tool(pfiles[name], args + ["ld=1", "ld2=1"])
        if "delta" in o:
            o["delta_star_dense"] = o.pop("delta_star")
            o["delta_star"] = o.pop("delta_star_refined")
            o["delta_ref_vs_star"] = np.array([np.linalg.norm(o["delta"] - o["delta_star"]) / np.linalg.norm(o["delta_star"])])
            o2 = run_tool(ulp_files[name], args)
            o["delta_ulp_sensitivity"] = np.array([np.linalg.norm(o2["delta"] - o["delta"]) / np.linalg.norm(o["delta"])])
        tag = f"{name}__solve_{ordering}_{lam:g}_{diag}"
        np.savez_compressed(os.path.join(OUT, tag + ".npz"), digest=digests[name], lam=lam, diag=diag, **o)
        print(tag, "cliques", len(o.get("clique_nfrontal", [])), "failed", int(o["failed"][0]), "ulp sensitivity",
              float(o.get("delta_ulp_sensitivity", [0])[0]), "|delta_ref - delta*| / |delta*|", float(o.get("delta_ref_vs_star", [0])[0]))
    if "--solves-only" in sys.argv:
        return
    for name, mode, ordering, extra in RUNS:
        o = run_tool(pfiles[name], [f"mode={mode}", f"ordering={ordering}"] + extra)
        tag = f"{name}__{mode}_{ordering}" + ("_ceres" if "ceres=1" in extra else "")
        np.savez_compressed(os.path.join(OUT, tag + ".npz"), digest=digests[name], **o)
        print(tag, "iterations", int(o["iterations"][0]), "final_error", float(o["final_error"][0]))


if __name__ == "__main__":
    main()
