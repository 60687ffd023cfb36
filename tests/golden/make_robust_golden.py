"""Golden fixtures for robust noise models (noiseModel::Robust), from the unmodified reference.

    make -C oracle -j8 ref && python tests/golden/make_robust_golden.py

Each case is one of the seeded problems of make_golden.py with an m-estimator wrapped around one bucket's noise."""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden  # noqa: E402

# name -> (base problem, bucket, m-estimator, parameter, ordering for solve / optimise)
CASES = {
    "pose2_300_huber": ("pose2_300", "between2", "huber", 1.345, "colamd"),
    "pose2_300_tukey": ("pose2_300", "between2", "tukey", 4.6851, "metis"),
    "pose3_120_cauchy": ("pose3_120", "between3", "cauchy", 1.0, "colamd"),
    "pose3_120_gemanmcclure": ("pose3_120", "between3", "gemanmcclure", 2.0, "metis"),
    "ba_30_500_huber": ("ba_30_500", "sfm", "huber", 1.0, "colamd"),
    "ba_30_500_fair": ("ba_30_500", "sfm", "fair", 1.3998, "metis"),
    "ba_30_500_welsch": ("ba_30_500", "sfm", "welsch", 2.9846, "colamd"),
    "ba_30_500_dcs": ("ba_30_500", "sfm", "dcs", 1.0, "colamd"),
    "ba_30_500_l2withdeadzone": ("ba_30_500", "sfm", "l2withdeadzone", 0.5, "colamd"),
}


def make_problem(name):
    base, bucket, loss, k, _ = CASES[name]
    p = make_golden.make_problem(base)
    p.set_robust(bucket, loss, k)
    return p


def main():
    tmp = tempfile.mkdtemp()
    for name, (_, bucket, loss, k, ordering) in CASES.items():
        p = make_problem(name)
        pf = os.path.join(tmp, name + ".gtb")
        p.save(pf)
        dig = make_golden.problem_digest(p)
        lin = make_golden.run_tool(pf, ["mode=linearize"])
        sol = make_golden.run_tool(pf, ["mode=solve", f"ordering={ordering}", "lambda=0.001", "diag=0"])
        lm = make_golden.run_tool(pf, ["mode=lm", f"ordering={ordering}"])
        out = {"digest": dig}
        out.update({"lin_" + k2: v for k2, v in lin.items()})
        out.update({"solve_" + k2: v for k2, v in sol.items()})
        out.update({"lm_" + k2: v for k2, v in lm.items()})
        np.savez_compressed(os.path.join(make_golden.OUT, f"robust__{name}.npz"), **out)
        print(name, "error0", float(lin["error0"][0]), "lm iterations", int(lm["iterations"][0]), "final", float(lm["final_error"][0]))


if __name__ == "__main__":
    main()
