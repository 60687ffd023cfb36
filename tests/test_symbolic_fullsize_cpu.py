"""Bit-exact clique layout of the Bayes tree at BASELINE.json's full sizes (C1 M3500-shape, C2 sphere2500-shape, C3
Ladybug-scale: 156,150 cliques, 855,390 keys): csrc/symbolic.cpp against the reference's own symbolic elimination
(EliminationTree -> JunctionTree -> BayesTree, inference/EliminateableFactorGraph-inst.h:125-146) run HERE through
oracle/_ref/ref_tool mode=symbolic on the same problem and the same (committed) ordering.  Integer work: exact."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_tool")

pytestmark = pytest.mark.skipif(not os.path.exists(TOOL), reason="oracle/_ref is built only where /root/reference exists")


@pytest.mark.parametrize("workload,ordering", [("c1", "metis"), ("c1", "colamd"), ("c2", "metis"), ("c2", "colamd"),
                                               ("c3", "metis")])
def test_fullsize_clique_layout_matches_the_reference(workload, ordering, tmp_path):
    sys.path.insert(0, ROOT)
    import bench
    from gtsam_petercdev_b200 import gtbio
    from gtsam_petercdev_b200.engine import Engine
    p = bench.make_problem(workload)
    keys = np.asarray(bench.ordering_for(p, workload, ordering), dtype=np.uint64)
    p.a["ordering"] = keys
    pf, out = str(tmp_path / "p.gtb"), str(tmp_path / "o.gtb")
    p.save(pf)
    subprocess.run([TOOL, pf, out, "mode=symbolic", "ordering=file"], check=True, timeout=900)
    g = gtbio.load(out)
    e = Engine(p, device=-1)
    e.set_ordering(keys)
    pl = e.plan()
    for k in ("clique_nfrontal", "clique_nkeys", "clique_keys", "clique_parent"):
        assert np.array_equal(pl[k], g[k]), k
    e.schedule_check()
