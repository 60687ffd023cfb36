import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, GOLDEN)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_sessionstart(session):
    # make sure the product library exists (nvcc cross-compiles without a GPU); no fallback if it cannot be built
    from gtsam_petercdev_b200 import build
    build.build()


def load_golden(tag):
    return dict(np.load(os.path.join(GOLDEN, tag + ".npz")))


_problems = {}


def golden_problem(name):
    """Regenerate a fixture's problem from its seed and check it against the digest stored with the oracle output."""
    import make_golden
    if name not in _problems:
        p = make_golden.make_problem(name)
        g = load_golden(name + "__linearize")
        assert make_golden.problem_digest(p) == str(g["digest"]), "synthetic generator drifted from the golden fixtures"
        _problems[name] = p
    return _problems[name]


def solve_tags():
    import make_golden
    return [(n, o, l, d, f"{n}__solve_{o}_{l:g}_{d}") for n, o, l, d in make_golden.SOLVES]


def run_tags():
    import make_golden
    return [(n, m, o, x, f"{n}__{m}_{o}" + ("_ceres" if "ceres=1" in x else "")) for n, m, o, x in make_golden.RUNS]


@pytest.fixture(scope="session")
def has_gpu():
    import torch
    return torch.cuda.is_available()
