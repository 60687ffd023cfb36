"""Known-answer fixture from the reference's own end-to-end test (tests/testGeneralSFMFactorB.cpp:44-63): BAL file
dubrovnik-3-7-pre -> default LevenbergMarquardtOptimizer -> graph.error == 0.0199833 (1e-5).

    make -C oracle -j8 ref && python tests/golden/make_bal_golden.py

Reads /root/reference/examples/Data/dubrovnik-3-7-pre.txt with gtsam_petercdev_b200.datasets.read_bal, runs the
unmodified reference (oracle/_ref/ref_tool mode=lm, COLAMD = the default ordering) and stores the parsed problem arrays
together with the reference's outputs, so the GPU test needs neither the file nor the reference."""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden  # noqa: E402
from gtsam_petercdev_b200 import datasets  # noqa: E402

BAL = "/root/reference/examples/Data/dubrovnik-3-7-pre.txt"
KNOWN_ANSWER = 0.0199833   # tests/testGeneralSFMFactorB.cpp:62


def main():
    p = datasets.read_bal(BAL)
    pf = os.path.join(tempfile.mkdtemp(), "bal.gtb")
    p.save(pf)
    o = make_golden.run_tool(pf, ["mode=lm", "ordering=colamd"])
    final = float(o["final_error"][0])
    assert abs(final - KNOWN_ANSWER) < 1e-5, final
    arrays = {"problem_" + k: v for k, v in p.a.items()}
    np.savez_compressed(os.path.join(make_golden.OUT, "bal_dubrovnik_3_7__lm_colamd.npz"), known_answer=KNOWN_ANSWER,
                        digest=make_golden.problem_digest(p), **arrays, **o)
    print("reference final error", final, "iterations", int(o["iterations"][0]))


if __name__ == "__main__":
    main()
