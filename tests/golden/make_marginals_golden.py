"""Golden fixtures for marginal covariances: the unmodified reference's
Marginals(graph, values).marginalCovariance(key) for EVERY variable of the small seeded problems.

    make -C oracle -j8 ref && python tests/golden/make_marginals_golden.py"""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden  # noqa: E402

PROBLEMS = ["pose2_300", "pose3_120", "ba_30_500"]


def main():
    tmp = tempfile.mkdtemp()
    for name in PROBLEMS:
        p = make_golden.make_problem(name)
        pf = os.path.join(tmp, name + ".gtb")
        p.save(pf)
        o = make_golden.run_tool(pf, ["mode=marginals"])
        np.savez_compressed(os.path.join(make_golden.OUT, f"marginals__{name}.npz"), digest=make_golden.problem_digest(p),
                            marginal_keys=o["marginal_keys"], marginal_dims=o["marginal_dims"], marginal_cov=o["marginal_cov"])
        print(name, len(o["marginal_keys"]), "variables")


if __name__ == "__main__":
    main()
