"""Marginal covariances from the device Bayes tree (marginals.cuh) against the reference's
Marginals::marginalCovariance for every variable (fixtures: tests/golden/make_marginals_golden.py).
Tolerance: 1e-8 of the block's norm (the reference itself inverts through a different elimination per query)."""
import numpy as np
import pytest

from conftest import golden_problem, load_golden
from gtsam_petercdev_b200.engine import Engine, EngineError

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,ordering", [("pose2_300", "colamd"), ("pose2_300", "metis"), ("pose3_120", "colamd"),
                                           ("pose3_120", "metis"), ("ba_30_500", "colamd"), ("ba_30_500", "metis"),
                                           ("ba_30_500", "schur")])
def test_marginal_covariances_match_the_reference(name, ordering):
    g = load_golden(f"marginals__{name}")
    tag = {"pose2_300": "pose2_300__solve_%s_0_0" % ordering if ordering == "colamd" else "pose2_300__solve_metis_0.001_0",
           "pose3_120": "pose3_120__solve_colamd_0_0" if ordering == "colamd" else "pose3_120__solve_metis_0.001_1",
           "ba_30_500": {"colamd": "ba_30_500__solve_colamd_1e-05_0", "metis": "ba_30_500__solve_metis_0.0001_1",
                         "schur": "ba_30_500__solve_schur_0.01_0"}[ordering]}[name]
    e = Engine(golden_problem(name), device=0)
    e.set_ordering(load_golden(tag)["ordering"])
    e.linearize()
    e.factorize(0.0)
    keys = g["marginal_keys"]
    covs = e.marginal_covariances(keys)
    o = 0
    worst = 0.0
    for k, d, C in zip(keys, g["marginal_dims"], covs):
        ref = g["marginal_cov"][o:o + d * d].reshape(d, d)
        o += d * d
        assert C.shape == (d, d)
        worst = max(worst, np.linalg.norm(C - ref) / np.linalg.norm(ref))
        assert np.allclose(C, C.T, rtol=0, atol=1e-12 * np.linalg.norm(ref))
    assert worst < 1e-8, worst


def test_marginals_api_errors():
    e = Engine(golden_problem("pose2_300"), device=0)
    e.set_ordering(load_golden("pose2_300__solve_colamd_0_0")["ordering"])
    e.linearize()
    with pytest.raises(EngineError):
        e.marginal_covariances(golden_problem("pose2_300").a["pose2_keys"][:1])   # not factorised yet
    e.factorize(0.0)
    with pytest.raises(EngineError):
        e.marginal_covariances(np.array([123456789], dtype=np.uint64))            # unknown key
    one = e.marginal_covariances(golden_problem("pose2_300").a["pose2_keys"][:2])
    assert len(one) == 2 and one[0].shape == (3, 3)
