"""The compiled drop-in (integration/DeviceOptimizers.{h,cpp}: subclasses of the reference's own LevenbergMarquardtOptimizer /
GaussNewtonOptimizer that run iterate() / linearize() / solve() on the device through the C ABI), exercised by
integration/dropin_test.cpp: the bodies of examples/Pose2SLAMExample_g2o.cpp, Pose3SLAMExample_g2o.cpp and SFMExample_bal.cpp
run with the stock optimizer and with the device optimizer on real NonlinearFactorGraph / Values objects, ordering from the
reference's own Ordering::Create (COLAMD and METIS).  The binary links the unmodified reference (oracle/_ref/libgtsam_ref.so)
and the product library; oracle/Makefile target `dropin` builds it (in the container that has /root/reference) and it travels
to the GPU box like ref_tool."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "dropin_test")
DATA = os.path.join(ROOT, "tests", "golden", "data")


@pytest.mark.gpu
def test_device_optimizers_match_the_stock_optimizers_on_the_reference_examples():
    assert os.path.exists(BIN), "oracle/_ref/dropin_test is missing: run `make -C oracle dropin` where /root/reference exists"
    r = subprocess.run([BIN, DATA], capture_output=True, text=True, timeout=600)
    print(r.stdout)
    print(r.stderr)
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "dropin_test.txt"), "w") as f:
        f.write(r.stdout)
    lines = [l for l in r.stdout.splitlines() if l.startswith(("PASS", "FAIL"))]
    assert r.returncode == 0 and lines and all(l.startswith("PASS") for l in lines), r.stdout + r.stderr
    assert len(lines) >= 14


def test_dropin_sources_declare_the_reference_seams():
    src = open(os.path.join(ROOT, "integration", "DeviceOptimizers.h")).read()
    for needle in ("class DeviceLevenbergMarquardtOptimizer : public LevenbergMarquardtOptimizer",
                   "class DeviceGaussNewtonOptimizer : public GaussNewtonOptimizer",
                   "GaussianFactorGraph::shared_ptr iterate() override", "GaussianFactorGraph::shared_ptr linearize() const override",
                   "VectorValues solve(const GaussianFactorGraph& gfg, const NonlinearOptimizerParams& params) const override"):
        assert needle in src
