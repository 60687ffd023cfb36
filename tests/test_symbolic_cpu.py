"""Bit-exact parity of the symbolic plan (clique layout of the Bayes tree) with the reference, on the committed
golden dumps of oracle/_ref/ref_tool (COLAMD, METIS, natural and Schur orderings).  Integer work: exact equality."""
import numpy as np
import pytest

from conftest import golden_problem, load_golden, solve_tags
from gtsam_petercdev_b200.engine import Engine


@pytest.mark.parametrize("name,ordering,lam,diag,tag", solve_tags())
def test_clique_layout_matches_reference(name, ordering, lam, diag, tag):
    g = load_golden(tag)
    e = Engine(golden_problem(name), device=-1)
    e.set_ordering(g["ordering"])
    pl = e.plan()
    for k in ("clique_nfrontal", "clique_nkeys", "clique_keys", "clique_parent"):
        assert np.array_equal(pl[k], g[k]), k
    # front layout consistency: every conditional block count adds up
    assert pl["n_scalars"] == int(e.var_dim.sum())


def test_builtin_orderings_match_reference():
    g = load_golden("ba_30_500__solve_schur_0.01_0")
    e = Engine(golden_problem("ba_30_500"), device=-1)
    e.set_ordering("schur")
    pl = e.plan()
    assert np.array_equal(pl["clique_keys"], g["clique_keys"])
    g = load_golden("pose2_300__solve_natural_0.01_1")
    e = Engine(golden_problem("pose2_300"), device=-1)
    e.set_ordering("natural")
    assert np.array_equal(e.plan()["clique_keys"], g["clique_keys"])
