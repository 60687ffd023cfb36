"""File -> native reader -> device LM against file -> reference reader -> reference LM (fixtures:
tests/golden/make_ingest_golden.py).  Final error within 1e-6 relative, same iteration count."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from gtsam_petercdev_b200 import datasets
from gtsam_petercdev_b200.engine import Engine

pytestmark = pytest.mark.gpu
import make_ingest_golden  # noqa: E402


@pytest.mark.parametrize("name,kind", [c for c in make_ingest_golden.FILES if c[0] != "pose3example-offdiagonal.txt"])
def test_loaded_problem_optimises_like_the_reference(name, kind):
    g = load_golden("ingest__" + name.replace(".txt", "").replace("-", "_"))
    path = os.path.join(GOLDEN, "data", name)
    p = datasets.read_bal_native(path) if kind == "bal" else datasets.read_g2o(path, kind == "g2o3")
    make_ingest_golden.with_prior(p, kind)
    e = Engine(p, device=0)
    e.set_ordering(g["lm_ordering"])
    assert abs(e.error() - g["lm_error0"][0]) <= 1e-9 * g["lm_error0"][0]
    r = e.lm_optimize()
    assert abs(r["final_error"] - g["lm_final_error"][0]) <= 1e-6 * g["lm_final_error"][0]
    assert r["iterations"] == int(g["lm_iterations"][0])
