"""Native dataset readers / writer (csrc/ingest.cpp behind the C ABI) against the reference's own
readG2o / SfmData::FromBalFile / writeG2o (fixtures: tests/golden/make_ingest_golden.py).  Integer and parsed-number
work is bit-exact; information matrices are compared with the reference's R^T R to 1e-12."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden
from gtsam_petercdev_b200 import datasets

DATA = os.path.join(GOLDEN, "data")
CASES = [("noisyToyGraph.txt", "g2o2"), ("pose2example.txt", "g2o2"), ("pose3example.txt", "g2o3"),
         ("pose3example-offdiagonal.txt", "g2o3"), ("dubrovnik-3-7-pre.txt", "bal")]


def _tag(name):
    return "ingest__" + name.replace(".txt", "").replace("-", "_")


def _read(name, kind):
    path = os.path.join(DATA, name)
    return datasets.read_bal_native(path) if kind == "bal" else datasets.read_g2o(path, kind == "g2o3")


@pytest.mark.parametrize("name,kind", CASES)
def test_reader_matches_the_reference(name, kind):
    compare_with_reference(_read(name, kind), load_golden(_tag(name)), kind)


def compare_with_reference(p, g, kind, chained=False):
    """p: problem from the native reader; g: arrays dumped by the reference's reader (ref_* keys).  chained: the file
    has no vertices and the poses were created by chaining odometry: the reference then holds products of cosines and
    sines (Rot2 keeps (c, s)), the flat tables hold theta = atan2(s, c), so cos(theta) is c up to a rounding."""
    if kind == "g2o2":
        assert np.array_equal(p.a["pose2_keys"], g["ref_pose2_keys"])
        assert np.array_equal(p.a["pose2"][:, :2], g["ref_pose2"][:, :2])
        cs = np.stack([np.cos(p.a["pose2"][:, 2]), np.sin(p.a["pose2"][:, 2])], axis=1)   # what the device keeps
        if chained:
            # the reference renormalises a product only when its norm is off by more than 1e-10 (Rot2.cpp:56-64)
            np.testing.assert_allclose(cs, g["ref_pose2_cs"], rtol=0, atol=1e-10)
            assert np.array_equal(p.a["pose2"][:, 2], g["ref_pose2"][:, 2])               # Rot2::theta() of the product
        else:
            assert np.array_equal(cs, g["ref_pose2_cs"])
        assert np.array_equal(p.a["between2_keys"], g["ref_between2_keys"])
        assert np.array_equal(p.a["between2_z"][:, :2], g["ref_between2_z"][:, :2])
        zcs = np.stack([np.cos(p.a["between2_z"][:, 2]), np.sin(p.a["between2_z"][:, 2])], axis=1)
        assert np.array_equal(zcs, g["ref_between2_cs"])
        assert np.all(p.a["between2_nk"] == 3)
        np.testing.assert_allclose(p.a["between2_noise"], g["ref_between2_info"], rtol=1e-12, atol=1e-12)
    elif kind == "g2o3":
        if "ref_pose3_keys" in g and len(g["ref_pose3_keys"]):
            assert np.array_equal(p.a["pose3_keys"], g["ref_pose3_keys"])
            assert np.array_equal(p.a["pose3"], g["ref_pose3"])        # Eigen's quaternion -> matrix, same operation order
        else:                                                          # edges only: load3D creates no values
            assert "pose3_keys" not in p.a or len(p.a["pose3_keys"]) == 0
        assert np.array_equal(p.a["between3_keys"], g["ref_between3_keys"])
        assert np.array_equal(p.a["between3_z"], g["ref_between3_z"])
        scale = np.abs(g["ref_between3_info"]).max()
        np.testing.assert_allclose(p.a["between3_noise"], g["ref_between3_info"], rtol=1e-12, atol=1e-12 * scale)
    else:
        assert np.array_equal(p.a["cam_keys"], g["ref_cam_keys"])
        assert np.array_equal(p.a["point_keys"], g["ref_point_keys"])
        assert np.array_equal(p.a["point"], g["ref_point"])            # floats widened to double
        assert np.array_equal(p.a["sfm_keys"], g["ref_sfm_keys"])      # track order
        assert np.array_equal(p.a["sfm_z"], g["ref_sfm_z"])            # v negated
        np.testing.assert_allclose(p.a["cam"], g["ref_cam"], rtol=0, atol=1e-14 * np.abs(g["ref_cam"]).max())
        assert np.array_equal(p.a["cam"][:, 12:], g["ref_cam"][:, 12:])


def test_bal_native_agrees_with_the_python_reader():
    a = datasets.read_bal_native(os.path.join(DATA, "dubrovnik-3-7-pre.txt"))
    b = datasets.read_bal(os.path.join(DATA, "dubrovnik-3-7-pre.txt"))
    for k in ("cam_keys", "point_keys", "sfm_keys", "sfm_z", "point"):
        assert np.array_equal(a.a[k], b.a[k]), k
    np.testing.assert_allclose(a.a["cam"], b.a["cam"], rtol=0, atol=1e-13)


def _tokens(text):
    rows = []
    for line in text.strip().splitlines():
        t = line.split()
        rows.append((t[0], [float(x) for x in t[1:]]))
    return rows


@pytest.mark.parametrize("name,kind", CASES[:4])
def test_writer_matches_the_reference_and_round_trips(name, kind, tmp_path):
    g = load_golden(_tag(name))
    p = _read(name, kind)
    out = str(tmp_path / "w.g2o")
    datasets.write_g2o(out, p)
    mine, ref = _tokens(open(out).read()), _tokens(bytes(g["ref_written"]).decode())
    assert [t for t, _ in mine] == [t for t, _ in ref]
    for (_, a), (_, b) in zip(mine, ref):
        np.testing.assert_allclose(a, b, rtol=2e-5, atol=2e-6)   # both print 6 significant digits
    q = datasets.read_g2o(out, kind == "g2o3")
    vt, bt = ("pose2", "between2") if kind == "g2o2" else ("pose3", "between3")
    assert np.array_equal(q.a[vt + "_keys"], p.a[vt + "_keys"]) and np.array_equal(q.a[bt + "_keys"], p.a[bt + "_keys"])
    np.testing.assert_allclose(q.a[vt], p.a[vt], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(q.a[bt + "_z"], p.a[bt + "_z"], rtol=1e-4, atol=1e-5)


def test_pure_odometry_file_creates_the_vertices(tmp_path):
    """load2D inserts Pose2() for an unknown first key and chains the measurement for the second
    (slam/dataset.cpp:541-545)."""
    f = tmp_path / "odo.g2o"
    f.write_text("EDGE_SE2 0 1 1.0 0.0 1.5707963267948966 1 0 0 1 0 1\nEDGE_SE2 1 2 1.0 0.0 0.0 1 0 0 1 0 1\n")
    p = datasets.read_g2o(str(f))
    assert list(p.a["pose2_keys"]) == [0, 1, 2]
    np.testing.assert_allclose(p.a["pose2"], [[0, 0, 0], [1, 0, np.pi / 2], [1, 1, np.pi / 2]], atol=1e-15)
    assert np.all(p.a["between2_nk"] == 3)   # Gaussian::Information; the engine's smart factory turns it into Unit


def test_robust_kernel_and_errors(tmp_path):
    p = datasets.read_g2o(os.path.join(DATA, "noisyToyGraph.txt"), kernel="huber")
    assert list(p.a["between2_robust"]) == [2.0, 1.345]              # Huber::Create(1.345), slam/dataset.cpp:283
    p = datasets.read_g2o(os.path.join(DATA, "noisyToyGraph.txt"), kernel="tukey")
    assert list(p.a["between2_robust"]) == [4.0, 4.6851]
    with pytest.raises(datasets.DatasetError, match="can not find file"):
        datasets.read_g2o(str(tmp_path / "missing.g2o"))
    bad = tmp_path / "bad.g2o"
    bad.write_text("VERTEX_SE2 0 0 0\n")
    with pytest.raises(datasets.DatasetError, match="malformed"):
        datasets.read_g2o(str(bad))
    lm = tmp_path / "lm.g2o"
    lm.write_text("VERTEX_SE2 0 0 0 0\nVERTEX_XY 1 2 3\n")
    with pytest.raises(datasets.DatasetError, match="landmark"):
        datasets.read_g2o(str(lm))
    zero = tmp_path / "zero.g2o"
    zero.write_text("EDGE_SE2 0 1 1 0 0 0 0 0 1 0 1\n")
    with pytest.raises(datasets.DatasetError, match="not G2O matrix order"):
        datasets.read_g2o(str(zero))
    with pytest.raises(datasets.DatasetError):
        datasets.read_bal_native(str(tmp_path / "missing.txt"))


def test_dataset_apply_fills_an_engine():
    """gtb_dataset_apply pushes the tables into an engine; with a host-only engine the symbolic plan can be built."""
    import ctypes as C
    from gtsam_petercdev_b200 import capi
    lib = capi.lib()
    h, e = C.c_void_p(), C.c_void_p()
    assert lib.gtb_dataset_read_g2o(os.fsencode(os.path.join(DATA, "pose2example.txt")), 0, 0, C.byref(h)) == capi.OK
    assert lib.gtb_create(-1, C.byref(e)) == capi.OK
    assert lib.gtb_dataset_apply(h, e) == capi.OK
    assert lib.gtb_finalize_problem(e) == capi.OK
    assert lib.gtb_set_ordering_builtin(e, 0) == capi.OK
    n = [C.c_int64() for _ in range(4)]
    assert lib.gtb_plan_sizes(e, *[C.byref(x) for x in n]) == capi.OK and n[0].value > 0
    lib.gtb_destroy(e)
    lib.gtb_dataset_free(h)
