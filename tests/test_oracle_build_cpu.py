"""The oracle is the unmodified reference compiled by oracle/Makefile (no CMake): it is only as good as that build.
`make -C oracle reftests` compiles 36 of the reference's OWN unit-test programs for the path (dense Cholesky,
Hessian / Jacobian factors, conditionals, noise models and robust losses, elimination / junction / Bayes trees, variable
index, ordering, Between / GeneralSFM factors, Pose2 / Pose3 / Cal3Bundler, dataset and BAL readers, NonlinearFactorGraph,
NonlinearOptimizer, Marginals, ISAM2) from where they lie under /root/reference against
oracle/_ref/libgtsam_ref.so and runs them.  All of them must pass."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libgtsam_ref.so")


@pytest.mark.skipif(not (os.path.isdir("/root/reference/gtsam") and os.path.exists(LIB)),
                    reason="needs the reference sources and the oracle build (this container only)")
def test_reference_unit_tests_pass_on_the_oracle_build():
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "-j8", "reftests"], capture_output=True, text=True,
                       timeout=1500)
    lines = [l for l in r.stdout.splitlines() if l.startswith(("PASS ", "FAIL "))]
    assert r.returncode == 0, "\n".join(lines) + r.stderr[-2000:]
    assert len(lines) == 36 and all(l.startswith("PASS ") and "no test failures" in l for l in lines), lines
    names = {l.split()[1].rstrip(":") for l in lines}
    for must in ("testCholesky", "testHessianFactor", "testNonlinearOptimizer", "testGeneralSFMFactorB", "testMarginals",
                 "testGaussianISAM2", "testDataset", "testSfmData", "testRobust", "testSymbolicBayesTree", "testVariableIndex"):
        assert must in names
