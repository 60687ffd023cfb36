"""Elimination orderings of the bench workloads, computed by the UNMODIFIED reference (Ordering::Colamd / Ordering::Metis
through oracle/_ref/ref_tool mode=ordering) and stored as permutations of the key-sorted variable ids.

    make -C oracle -j8 ref && python tests/golden/make_orderings.py [ladybug_shape m3500_shape sphere_shape venice_shape]

The ordering is an INPUT of the hot path (the reference computes it once in the optimiser's constructor,
nonlinear/LevenbergMarquardtParams.h:112-117), so the GPU box gets it from these files."""
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden  # noqa: E402
from gtsam_petercdev_b200 import datasets  # noqa: E402

SHAPES = {"ladybug_shape": datasets.ladybug_shape, "m3500_shape": datasets.m3500_shape,
          "sphere_shape": datasets.sphere_shape, "venice_shape": datasets.venice_shape}


def main(names):
    tmp = tempfile.mkdtemp()
    for name in names:
        p = SHAPES[name]()
        keys = p.variables()[0]
        pf = os.path.join(tmp, name + ".gtb")
        p.save(pf)
        out = {}
        for o in ("colamd", "metis"):
            t0 = time.time()
            r = make_golden.run_tool(pf, ["mode=ordering", f"ordering={o}"])
            perm = np.searchsorted(keys, r["ordering"]).astype(np.int32)
            assert np.array_equal(keys[perm], r["ordering"]) and len(np.unique(perm)) == len(keys)
            out[o] = perm
            print(name, o, len(perm), "variables", f"{time.time() - t0:.1f} s")
        np.savez_compressed(os.path.join(make_golden.OUT, name + "_orderings.npz"), **out)


if __name__ == "__main__":
    main(sys.argv[1:] or ["ladybug_shape", "m3500_shape", "sphere_shape"])
