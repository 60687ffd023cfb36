"""The algorithmic work figures bench.py divides by the measured time (roofline.achieved) are computed by the library from
the plan (gtb_plan_stats).  Here they are restated in numpy from the public plan arrays and the problem sizes, on the
golden problems and on the full-size C3 workload: elimination flops per clique F^3/3 + F^2 (S+1) + F (S+1)(S+2)
(POTRF + TRSM + SYRK of base/cholesky.cpp:108-159 on an (F+S+1)-wide front), back-substitution bytes, and the
linearize bytes of SURVEY 8(d) (C3: 161.5 MB)."""
import numpy as np
import pytest

from conftest import golden_problem, load_golden, solve_tags
from gtsam_petercdev_b200.engine import Engine


def clique_dims(e, pl):
    dim_of_key = dict(zip(e.var_keys.tolist(), e.var_dim.tolist()))
    F, S, off = [], [], 0
    for nf, nk in zip(pl["clique_nfrontal"].tolist(), pl["clique_nkeys"].tolist()):
        d = [dim_of_key[k] for k in pl["clique_keys"][off:off + nk].tolist()]
        F.append(sum(d[:nf]))
        S.append(sum(d[nf:]))
        off += nk
    return np.array(F, dtype=np.float64), np.array(S, dtype=np.float64)


def check(e):
    pl = e.plan()
    st = pl["stats"]
    F, S = clique_dims(e, pl)
    s1 = S + 1
    flops = F ** 3 / 3 + F * F * s1 + F * s1 * (s1 + 1)
    assert st["flops"] == pytest.approx(flops.sum(), rel=1e-12)
    assert st["max_front"] == (F + S + 1).max()
    assert st["n_small"] + st["n_large"] == len(F)
    assert st["bytes_backsolve"] == pytest.approx((8 * (0.5 * F * (F + 1) + F * s1 + F)).sum(), rel=1e-12)
    assert 0 <= st["flops_large"] <= st["flops"] * (1 + 1e-12)
    assert st["flops_syrk_schur"] + st["flops_syrk_frontal"] <= st["flops_large"] * (1 + 1e-9)
    assert st["bytes_thin"] >= 0 and st["bytes_point"] >= 0 and st["bytes_dense_small"] >= 0
    return st, F, S


@pytest.mark.parametrize("name,ordering,lam,diag,tag", solve_tags())
def test_plan_statistics_restated(name, ordering, lam, diag, tag):
    e = Engine(golden_problem(name), device=-1)
    e.set_ordering(load_golden(tag)["ordering"])
    check(e)


def test_c3_figures_of_the_survey():
    import bench
    p = bench.make_problem("c3")
    e = Engine(p, device=-1)
    e.set_ordering(bench.ordering_for(p, "c3", "metis"))
    st, F, S = check(e)
    # cameras 120 B, points 24 B, per projection 8 B of key indices + 16 B measurement + 208 B of [A|b]
    n_cam, n_pt = (e.var_type == 2).sum(), (e.var_type == 3).sum()
    n_obs = int(p.factor_count(2)) if hasattr(p, "factor_count") else 679032
    want = n_cam * 120 + n_pt * 24 + n_obs * (8 + 16 + 208)
    assert n_cam == 1723 and n_pt == 156000
    assert st["bytes_linearize"] == pytest.approx(want, rel=2e-3)          # + two priors
    assert 161.0e6 < st["bytes_linearize"] < 162.0e6
    # point cliques: 208 B of factor rows read per projection, 32 B written per front column
    is_pt = (F == 3) & (S % 9 == 0) & (S > 0) & (S <= 9 * 64)
    assert st["bytes_point"] <= (208 * (S[is_pt] / 9) + 32 * (F[is_pt] + S[is_pt] + 1)).sum() * (1 + 1e-12)
    assert st["bytes_point"] > 0.95 * 679032 * 208
