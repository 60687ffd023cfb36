"""Golden fixtures for the native dataset readers / writer.

    make -C oracle -j8 ref && python tests/golden/make_ingest_golden.py

For every small data file committed under tests/golden/data/ (copies of the reference's examples/Data files) the
UNMODIFIED reference reads it with its own readG2o / SfmData::FromBalFile (oracle/_ref/ref_tool mode=readg2o|readbal) and
dumps variables, measurements and the information matrix of every noise model; it also writes the graph back with its
own writeG2o (mode=writeg2o).  Then the reference optimises the loaded problem (prior on the first pose as in
examples/Pose2SLAMExample_g2o.cpp:65-67 and Pose3SLAMExample_g2o.cpp:42-48) so that the GPU test can check the whole
chain file -> native reader -> device LM against file -> reference reader -> reference LM."""
import os
import subprocess
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden  # noqa: E402
from gtsam_petercdev_b200 import datasets, gtbio  # noqa: E402

DATA = os.path.join(make_golden.OUT, "data")
FILES = [("noisyToyGraph.txt", "g2o2"), ("pose2example.txt", "g2o2"), ("pose3example.txt", "g2o3"),
         ("pose3example-offdiagonal.txt", "g2o3"), ("dubrovnik-3-7-pre.txt", "bal")]


def with_prior(p, kind):
    """The prior the reference's examples add before optimising."""
    if kind == "g2o2":
        k0 = p.a["pose2_keys"][:1]
        p.set_factors("prior2", k0, p.a["pose2"][:1], 5, np.array([[1e-6, 1e-6, 1e-8]]))
    elif kind == "g2o3":
        k0 = p.a["pose3_keys"][:1]
        p.set_factors("prior3", k0, p.a["pose3"][:1], 5, np.array([[1e-6, 1e-6, 1e-6, 1e-4, 1e-4, 1e-4]]))
    return p


def main():
    for name, kind in FILES:
        path = os.path.join(DATA, name)
        out = tempfile.mktemp(suffix=".gtb")
        args = ["mode=readbal"] if kind == "bal" else ["mode=readg2o", "is3d=%d" % (kind == "g2o3")]
        subprocess.run([make_golden.TOOL, path, out] + args, check=True)
        ref = gtbio.load(out)
        arrays = {"ref_" + k: v for k, v in ref.items()}
        if kind != "bal":
            wr = tempfile.mktemp(suffix=".g2o")
            subprocess.run([make_golden.TOOL, path, wr, "mode=writeg2o", "is3d=%d" % (kind == "g2o3")], check=True)
            arrays["ref_written"] = np.frombuffer(open(wr, "rb").read(), dtype=np.uint8)
        # the reference optimises the problem it loaded (arrays identical to the native reader's: checked in the CPU test)
        p = datasets.read_bal_native(path) if kind == "bal" else datasets.read_g2o(path, kind == "g2o3")
        with_prior(p, kind)
        pf = tempfile.mktemp(suffix=".gtb")
        p.save(pf)
        o = make_golden.run_tool(pf, ["mode=lm", "ordering=colamd"])
        tag = "ingest__" + name.replace(".txt", "").replace("-", "_")
        np.savez_compressed(os.path.join(make_golden.OUT, tag + ".npz"), kind=kind, **arrays,
                            **{"lm_" + k: v for k, v in o.items()})
        print(tag, "reference error %.6g -> %.6g in %d iterations" % (o["error0"][0], o["final_error"][0], int(o["iterations"][0])))


if __name__ == "__main__":
    main()
