"""C-ABI surface and host-side behaviour that needs no GPU."""
import ctypes as C

import numpy as np
import pytest

from gtsam_petercdev_b200 import capi, datasets
from gtsam_petercdev_b200.engine import Engine, EngineError
from gtsam_petercdev_b200.problem import Problem


def test_library_exports_every_header_symbol():
    lib = capi.lib()
    names = capi.header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gtsam_b200.h but not exported"
    assert set(names) == set(capi._SIGS), "ctypes signature table out of sync with the header"
    assert b"sm_100a" in lib.gtb_version()


def test_host_only_engine_refuses_numeric_calls():
    p = datasets.m3500_shape(200, 10, seed=1)
    e = Engine(p, device=-1)
    e.set_ordering("natural")
    assert e.plan()["stats"]["levels"] >= 1
    for call in (e.linearize, e.backsolve, e.retract, lambda: e.error(0), lambda: e.factorize(0.0)):
        with pytest.raises(EngineError) as ei:
            call()
        assert ei.value.status in (capi.ERR_NO_DEVICE, capi.ERR_INVALID)
    with pytest.raises(EngineError) as ei:
        e.linearize()
    assert ei.value.status == capi.ERR_NO_DEVICE


def test_missing_key_and_wrong_type_are_rejected():
    p = datasets.m3500_shape(200, 5, seed=2)
    a = dict(p.a)
    a["between2_keys"] = a["between2_keys"].copy()
    a["between2_keys"][3, 1] = 10 ** 6
    with pytest.raises(EngineError) as ei:
        Engine(Problem(a), device=-1)
    assert ei.value.status == capi.ERR_KEY_MISSING


def test_constrained_noise_is_rejected():
    p = datasets.m3500_shape(200, 5, seed=2)
    a = dict(p.a)
    a["prior2_nk"] = np.array([2], dtype=np.int32)
    a["prior2_noise"] = np.array([[1e-9, 1.0, 1.0, 0, 0, 0, 0, 0, 0]])
    with pytest.raises(EngineError) as ei:
        Engine(Problem(a), device=-1)
    assert ei.value.status == capi.ERR_NOISE_MODEL


def test_bad_ordering_is_rejected():
    p = datasets.m3500_shape(200, 5, seed=2)
    e = Engine(p, device=-1)
    keys = p.a["pose2_keys"]
    with pytest.raises(EngineError):
        e.set_ordering(keys[:-1])
    with pytest.raises(EngineError):
        e.set_ordering(np.concatenate([keys[:-1], keys[:1]]))
    with pytest.raises(EngineError):
        e.set_ordering("schur")


def test_unconnected_variable_is_inconsistent():
    p = datasets.m3500_shape(200, 5, seed=2)
    a = dict(p.a)
    a["pose2_keys"] = np.concatenate([a["pose2_keys"], np.array([999], dtype=np.uint64)])
    a["pose2"] = np.concatenate([a["pose2"], np.zeros((1, 3))])
    e = Engine(Problem(a), device=-1)
    with pytest.raises(EngineError) as ei:
        e.set_ordering("natural")
    assert ei.value.status == capi.ERR_INCONSISTENT


def test_isam2_layer_host_side():
    """gtb_isam2_*: argument checks and staging work without a device; the update itself needs one and says so."""
    import numpy as np
    from gtsam_petercdev_b200.engine import EngineError
    from gtsam_petercdev_b200.isam2 import ISAM2
    with pytest.raises(EngineError):
        ISAM2(relinearize_threshold=-1.0)
    with pytest.raises(EngineError):
        ISAM2(relinearize_skip=0)
    s = ISAM2(0.01, 1, device=-1)
    x = lambda i: (ord("x") << 56) | i
    s.add_variables("pose2", [x(0)], [[0.0, 0.0, 0.0]])
    with pytest.raises(EngineError, match="already exists"):
        s.add_variables("pose2", [x(0)], [[1.0, 0.0, 0.0]])
    s.add_factors("prior2", [x(0)], [[0.0, 0.0, 0.0]], 2, [[1e-4, 1e-4, 1e-4]])
    assert s.count("pose2") == 0                      # staged, not committed
    with pytest.raises(EngineError) as ei:
        s.update()
    assert ei.value.status == capi.ERR_NO_DEVICE
    assert s.count("pose2") == 1                      # the additions stay committed, as in the reference
    with pytest.raises(EngineError):
        s.calculate_estimate("pose2")
    # a factor on an unknown key is refused by the engine at the update
    s.add_factors("between2", [[x(0), x(7)]], [[1.0, 0.0, 0.0]])
    with pytest.raises(EngineError) as ei:
        s.update()
    assert ei.value.status == capi.ERR_KEY_MISSING
    s.close()


def test_isam2_layer_plans_every_update_with_a_valid_ordering():
    """The ordering an ISAM2 update eliminates with -- the library's nested dissection every 8th update, in between the last one
    with the new variables appended, or the caller's order with the rest appended -- is a permutation of the variables at every
    update.  Checked on the host: the update stops at the first numeric call (no device), after the plan was built."""
    import ctypes as C
    from gtsam_petercdev_b200.engine import EngineError, _ptr
    from gtsam_petercdev_b200.isam2 import ISAM2, X
    lib = capi.lib()

    def frontal_keys(s):
        eng = lib.gtb_isam2_engine(s.h)
        nc, nk = C.c_int64(), C.c_int64()
        assert lib.gtb_plan_sizes(eng, C.byref(nc), C.byref(nk), None, None) == capi.OK
        nf, nkeys = np.zeros(nc.value, dtype=np.int64), np.zeros(nc.value, dtype=np.int64)
        keys = np.zeros(nk.value, dtype=np.uint64)
        assert lib.gtb_plan_get(eng, _ptr(nf), _ptr(nkeys), _ptr(keys), None) == capi.OK
        starts = np.cumsum(nkeys) - nkeys
        return np.concatenate([keys[a:a + f] for a, f in zip(starts, nf)])

    for given in (False, True):
        s = ISAM2(0.01, 1, device=-1)
        if given:
            s.set_ordering([X(i) for i in (5, 3, 40, 1, 0, 2, 4)])     # names a key that never exists: skipped
        s.add_variables("pose2", [X(0)], [[0.0, 0.0, 0.0]])
        s.add_factors("prior2", [X(0)], [[0.0, 0.0, 0.0]])
        for i in range(1, 21):
            with pytest.raises(EngineError) as ei:
                s.update()
            assert ei.value.status == capi.ERR_NO_DEVICE
            got = frontal_keys(s)
            assert sorted(got.tolist()) == [X(k) for k in range(i)], (given, i)
            s.add_variables("pose2", [X(i)], [[float(i), 0.0, 0.0]])
            s.add_factors("between2", [[X(i - 1), X(i)]], [[1.0, 0.0, 0.0]])
            if i % 5 == 0:
                s.add_factors("between2", [[X(0), X(i)]], [[float(i), 0.0, 0.0]])   # a loop closure
        s.close()
