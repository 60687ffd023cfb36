"""Native readers against the reference's readers over the reference's OWN data directory (examples/Data), beyond the
five files committed as fixtures: every g2o / TORO / BAL file there that holds only the factor types of the path.
Needs /root/reference and oracle/_ref (this container); same bit-exact comparison as tests/test_ingest_cpu.py."""
import os
import subprocess
import sys

import numpy as np
import pytest

from test_ingest_cpu import compare_with_reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = "/root/reference/examples/Data"
TOOL = os.path.join(ROOT, "oracle", "_ref", "ref_tool")
# (only files the reference SHIPS: "*-rewritten.txt" by-products of the reference's own testDataset are not inputs)
FILES = [("optimizedNoisyToyGraph.txt", "g2o2"), ("orientationsNoisyToyGraph.txt", "g2o2"),
         ("w20000.txt", "g2o2"),
         ("Klaus3.g2o", "g2o3"), ("toyExample.g2o", "g2o3"), ("pose3Localizationexample.txt", "g2o3"),
         ("pose3example-grid.txt", "g2o3"),
         ("pose3example-offdiagonal-rewritten.txt", "g2o3"), ("simpleGraph10gradIter.txt", "g2o3"),
         ("sphere2500.txt", "g2o3"), ("sphere2500_groundtruth.txt", "g2o3"),
         ("dubrovnik-3-7-18-pre.txt", "bal"), ("dubrovnik-1-1-pre.txt", "bal")]

pytestmark = pytest.mark.skipif(not (os.path.isdir(DATA) and os.path.exists(TOOL)),
                                reason="needs the reference's data directory and the oracle build (this container only)")


@pytest.mark.parametrize("name,kind", FILES)
def test_reader_matches_the_reference_on_its_own_data(name, kind, tmp_path):
    sys.path.insert(0, ROOT)
    from gtsam_petercdev_b200 import datasets, gtbio
    path = os.path.join(DATA, name)
    out = str(tmp_path / "ref.gtb")
    args = ["mode=readbal"] if kind == "bal" else ["mode=readg2o", "is3d=%d" % (kind == "g2o3")]
    subprocess.run([TOOL, path, out] + args, check=True, timeout=600)
    g = {"ref_" + k: v for k, v in gtbio.load(out).items()}
    p = datasets.read_bal_native(path) if kind == "bal" else datasets.read_g2o(path, kind == "g2o3")
    compare_with_reference(p, g, kind, chained=name == "w20000.txt")


def test_covariance_free_toro_file_is_rejected(tmp_path):
    """sphere_smallnoise.graph has EDGE3 lines without an information matrix.  The reference's 3D reader does not
    notice: it returns ONE factor with an all-zero information matrix for the 8,647 edges of the file.  The native
    reader refuses the file instead of handing a singular problem to the device."""
    sys.path.insert(0, ROOT)
    from gtsam_petercdev_b200 import datasets, gtbio
    path = os.path.join(DATA, "sphere_smallnoise.graph")
    out = str(tmp_path / "ref.gtb")
    subprocess.run([TOOL, path, out, "mode=readg2o", "is3d=1"], check=True, timeout=600)
    g = gtbio.load(out)
    assert g["between3_keys"].shape[0] == 1 and not g["between3_info"].any()
    with pytest.raises(datasets.DatasetError):
        datasets.read_g2o(path, True)


WRITE = [(n, k) for n, k in FILES if k != "bal" and n not in ("w20000.txt", "sphere2500.txt", "sphere2500_groundtruth.txt")]


@pytest.mark.parametrize("name,kind", WRITE + [("w20000.txt", "g2o2")])
def test_writer_matches_the_reference_on_its_own_data(name, kind, tmp_path):
    """gtb_write_g2o against the reference's writeG2o (slam/dataset.cpp:636-760), record for record; both print 6
    significant digits."""
    sys.path.insert(0, ROOT)
    from gtsam_petercdev_b200 import datasets
    from test_ingest_cpu import _tokens
    path = os.path.join(DATA, name)
    ref_out, my_out = str(tmp_path / "ref.g2o"), str(tmp_path / "mine.g2o")
    subprocess.run([TOOL, path, ref_out, "mode=writeg2o", "is3d=%d" % (kind == "g2o3")], check=True, timeout=600)
    datasets.write_g2o(my_out, datasets.read_g2o(path, kind == "g2o3"))
    mine, ref = _tokens(open(my_out).read()), _tokens(open(ref_out).read())
    assert [t for t, _ in mine] == [t for t, _ in ref]
    for (_, a), (_, b) in zip(mine, ref):
        np.testing.assert_allclose(a, b, rtol=2e-5, atol=2e-6)
