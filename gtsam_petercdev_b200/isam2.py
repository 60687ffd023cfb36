"""Host-side handle on the ISAM2 layer of the C ABI (include/gtsam_b200.h, gtb_isam2_*): a typed wrapper, no numerics here.

Names follow gtsam::ISAM2 (nonlinear/ISAM2.h): update(newFactors, newTheta) -> result, calculateEstimate().  The factors and
values of an update are given per bucket / variable type in the row layouts of gtb_set_factors / gtb_set_variables.
"""
import ctypes as C

import numpy as np

from . import capi
from .engine import EngineError, IndeterminantLinearSystemException, _ptr
from .problem import BUCKETS, BUCKET_DIM, VTYPES

_STATE_DEV = [4, 12, 17, 3]


class ISAM2:
    def __init__(self, relinearize_threshold=0.1, relinearize_skip=10, device=0):
        """Defaults are ISAM2Params' (nonlinear/ISAM2Params.h); examples/ISAM2_City10000.cpp:66-68 uses 0.01 and 1."""
        self.lib = capi.lib()
        h = C.c_void_p()
        st = self.lib.gtb_isam2_create(int(device), float(relinearize_threshold), int(relinearize_skip), C.byref(h))
        if st != capi.OK:
            raise EngineError(st, "gtb_isam2_create: bad parameters")
        self.h = h

    def _ck(self, st, key=None):
        if st == capi.OK:
            return
        msg = self.lib.gtb_isam2_last_error(self.h).decode()
        if st == capi.ERR_INDETERMINATE:
            raise IndeterminantLinearSystemException(st, msg, key.value if key is not None else 0)
        raise EngineError(st, msg)

    def close(self):
        if getattr(self, "h", None):
            self.lib.gtb_isam2_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_variables(self, vtype, keys, state):
        t = VTYPES.index(vtype) if isinstance(vtype, str) else vtype
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        state = np.ascontiguousarray(state, dtype=np.float64).reshape(len(keys), -1)
        self._ck(self.lib.gtb_isam2_add_variables(self.h, t, len(keys), _ptr(keys), _ptr(state)))

    def add_factors(self, bucket, keys, z, noise_kind=None, noise=None):
        b = BUCKETS.index(bucket) if isinstance(bucket, str) else bucket
        z = np.ascontiguousarray(z, dtype=np.float64)
        z = z.reshape(-1, z.shape[-1]) if z.ndim > 1 else z.reshape(1, -1)
        keys = np.ascontiguousarray(keys, dtype=np.uint64).reshape(len(z), -1)
        nk = nz = None
        if noise_kind is not None:
            nk = np.ascontiguousarray(np.broadcast_to(np.asarray(noise_kind, dtype=np.int32), (len(z),)))
            nz = np.zeros((len(z), BUCKET_DIM[b] ** 2))
            noise = np.asarray(noise, dtype=np.float64).reshape(len(z), -1)
            nz[:, :noise.shape[1]] = noise          # left-packed, as gtb_set_factors takes it
        self._ck(self.lib.gtb_isam2_add_factors(self.h, b, len(z), _ptr(keys), _ptr(z), _ptr(nk), _ptr(nz)))

    def set_ordering(self, keys):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        self._ck(self.lib.gtb_isam2_set_ordering(self.h, len(keys), _ptr(keys)))

    def update(self):
        """ISAM2::update on what was added since the last call.  Returns the ISAM2Result fields as a dict."""
        r = capi.ISAM2Result()
        key = C.c_uint64(0)
        self._ck(self.lib.gtb_isam2_update(self.h, C.byref(r), C.byref(key)), key)
        return {f: getattr(r, f) for f, _ in capi.ISAM2Result._fields_}

    def count(self, vtype):
        t = VTYPES.index(vtype) if isinstance(vtype, str) else vtype
        return int(self.lib.gtb_isam2_count(self.h, t))

    def calculate_estimate(self, vtype):
        """theta (+) delta of every variable of one type, rows in the order they were added (device row layout: POSE2 rows
        are (cos, sin, x, y))."""
        t = VTYPES.index(vtype) if isinstance(vtype, str) else vtype
        out = np.zeros((self.count(t), _STATE_DEV[t]))
        self._ck(self.lib.gtb_isam2_estimate(self.h, t, _ptr(out)))
        return out


# ---- the City10000 schedule (examples/ISAM2_City10000.cpp:62-134), shared by tests/test_isam2_gpu.py and bench.py --------
DIAG_SIGMAS = 2   # GTB_NOISE_DIAG_SIGMAS


def X(i):
    return (ord("x") << 56) | int(i)   # gtsam::Symbol('x', i)


def read_city10000(path):
    """Lines "EDGE2 <from> <n> <to> <n> <hypotheses> {x y theta}...": (from, to, x, y, theta) of the first hypothesis."""
    rows = []
    for line in open(path):
        f = line.split()
        if len(f) >= 9 and f[0] == "EDGE2":
            rows.append((int(f[1]), int(f[3]), float(f[6]), float(f[7]), float(f[8])))
    return rows


def pose2_compose(p, z):
    """Pose2 * Pose2 (geometry/Pose2.h:147-151) for an estimate row (cos, sin, x, y) and a measurement (x, y, theta)."""
    c, s, x, y = p
    cz, sz = np.cos(z[2]), np.sin(z[2])
    return [x + c * z[0] - s * z[1], y + s * z[0] + c * z[1], np.arctan2(s * cz + c * sz, c * cz - s * sz)]


def drive_city10000(isam, rows, on_update=None):
    """Prior 1e-4 on pose 0 and one update; then one update per line: an odometry line (to = from + 1) adds the new pose,
    initialised by composing the current estimate of `from` with the measurement, and a BetweenFactor with sigmas
    (1/30, 1/30, 1/100); any other line is a loop closure with sigmas 10.  The estimate is read back after every update, as
    the example does.  Returns (estimate rows, pose -> row, the result dict of every update)."""
    isam.add_variables("pose2", [X(0)], [[0.0, 0.0, 0.0]])
    isam.add_factors("prior2", [X(0)], [[0.0, 0.0, 0.0]], DIAG_SIGMAS, [[1e-4, 1e-4, 1e-4]])
    isam.update()
    est = isam.calculate_estimate("pose2")
    row_of = {0: 0}
    results = []
    for u, (a, b, zx, zy, zt) in enumerate(rows, start=1):
        z = [zx, zy, zt]
        if b == a + 1:
            row_of[b] = len(row_of)
            isam.add_variables("pose2", [X(b)], [pose2_compose(est[row_of[a]], z)])
            isam.add_factors("between2", [[X(a), X(b)]], [z], DIAG_SIGMAS, [[1.0 / 30.0, 1.0 / 30.0, 1.0 / 100.0]])
        else:
            isam.add_factors("between2", [[X(a), X(b)]], [z], DIAG_SIGMAS, [[10.0, 10.0, 10.0]])
        results.append(isam.update())
        est = isam.calculate_estimate("pose2")
        if on_update:
            on_update(u, est, row_of)
    return est, row_of, results
