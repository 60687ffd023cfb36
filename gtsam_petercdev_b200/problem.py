"""Flat (struct-of-arrays) description of a nonlinear least-squares problem.

This is the host-side interchange format between the reference-facing API
(NonlinearFactorGraph + Values, see ``api.py``), the CUDA engine (C-ABI in
``include/gtsam_b200.h``) and the test oracle (``oracle/ref_tool.cpp`` reads the
same arrays from a GTB1 file and builds the reference's own objects).

Variable arrays (all float64; keys uint64 as gtsam::Key, inference/Symbol.cpp:29-48):
  pose2_keys[n], pose2[n,3]   (x, y, theta)
  pose3_keys[n], pose3[n,12]  (R row-major 9, t 3)
  cam_keys[n],   cam[n,17]    (R 9, t 3, f, k1, k2, u0, v0)  PinholeCamera<Cal3Bundler>
  point_keys[n], point[n,3]

Factor buckets  <b>_keys[m,(2|1)] uint64, <b>_z[m,Z], <b>_nk[m] int32, <b>_noise[m,d*d]:
  0 between2  BetweenFactor<Pose2>       Z=3  d=3   slam/BetweenFactor.h:111
  1 between3  BetweenFactor<Pose3>       Z=12 d=6
  2 sfm       GeneralSFMFactor<Cal3Bundler camera, Point3>  Z=2 d=2  slam/GeneralSFMFactor.h:141
  3 prior2    PriorFactor<Pose2>         Z=3  d=3   nonlinear/PriorFactor.h:98
  4 prior3    PriorFactor<Pose3>         Z=12 d=6
  5 priorcam  PriorFactor<PinholeCamera<Cal3Bundler>>  Z=17 d=9
  6 priorpt   PriorFactor<Point3>        Z=3  d=3
Noise kinds as GIVEN by the user (nk): 0 Unit, 1 Isotropic sigma, 2 Diagonal sigmas,
  3 Gaussian::Information(matrix), 4 Gaussian::SqrtInformation(R), 5 Diagonal::Variances,
  all with the reference's smart=true downgrade (linear/NoiseModel.cpp:84-132,284-315,625-634).
Optional graph_order[M,2] int64 = (bucket, index) of every factor in graph order; default is
bucket-major in the order above.
"""
import numpy as np

BUCKETS = ["between2", "between3", "sfm", "prior2", "prior3", "priorcam", "priorpt"]
# m-estimators of linear/LossFunctions.h, ids as GTB_LOSS_* in include/gtsam_b200.h
LOSSES = ["none", "fair", "huber", "cauchy", "tukey", "welsch", "gemanmcclure", "dcs", "l2withdeadzone"]
BUCKET_Z = [3, 12, 2, 3, 12, 17, 3]
BUCKET_DIM = [3, 6, 2, 3, 6, 9, 3]          # residual dimension m
BUCKET_NKEYS = [2, 2, 2, 1, 1, 1, 1]
# variable type of every key slot of a bucket: 0 pose2, 1 pose3, 2 cam, 3 point
BUCKET_VTYPES = [(0, 0), (1, 1), (2, 3), (0,), (1,), (2,), (3,)]
VTYPES = ["pose2", "pose3", "cam", "point"]
VTYPE_STATE = [3, 12, 17, 3]
VTYPE_DIM = [3, 6, 9, 3]                    # tangent dimension

# resolved (device) noise kinds
NK_UNIT, NK_ISO, NK_DIAG, NK_FULL = 0, 1, 2, 3


def symbol(c, j):
    """gtsam::Symbol(c, j).key()  (inference/Symbol.cpp:29-48)."""
    return (np.uint64(ord(c)) << np.uint64(56)) | np.asarray(j, dtype=np.uint64)


class InvalidNoiseModel(ValueError):
    """Constrained (sigma < 1e-8) or otherwise unsupported noise model: rejected, no CPU fallback."""


def _chol_upper(M):
    """R with R^T R = M, as Eigen::LLT<Matrix>(M).matrixU() (linear/NoiseModel.cpp:108-110)."""
    return np.linalg.cholesky(M).T


def resolve_noise(kind, data, d):
    """Apply the reference's smart factory rules to one model.

    Returns (resolved_kind, R) with R a d*d row-major whitening matrix whose meaning depends
    on the kind: unit -> ignored; iso -> R[0] = 1/sigma; diag -> R[0:d] = 1/sigmas;
    full -> upper-triangular sqrt information.
    """
    out = np.zeros(d * d)
    data = np.asarray(data, dtype=np.float64)

    def iso_sigma(sigma):
        if abs(sigma - 1.0) < 1e-9:
            return NK_UNIT, out
        out[0] = 1.0 / sigma
        return NK_ISO, out

    def iso_variance(var):
        if abs(var - 1.0) < 1e-9:
            return NK_UNIT, out
        out[0] = 1.0 / np.sqrt(var)
        return NK_ISO, out

    def diag_sigmas(s):
        if np.any(s < 1e-8):
            raise InvalidNoiseModel("constrained noise model (sigma < 1e-8) is not supported on device")
        if np.all(s == s[0]):
            return iso_sigma(float(s[0]))
        out[:d] = 1.0 / s
        return NK_DIAG, out

    def diag_variances(v):
        if np.all(v == v[0]):
            return iso_variance(float(v[0]))
        out[:d] = 1.0 / np.sqrt(v)
        return NK_DIAG, out

    def is_diag(M):
        return not np.any(np.abs(np.triu(M, 1)) > 1e-9)

    if kind == 0:
        return NK_UNIT, out
    if kind == 1:
        return iso_sigma(float(data[0]))
    if kind == 2:
        return diag_sigmas(data[:d].copy())
    if kind == 5:
        return diag_variances(data[:d].copy())
    M = data[: d * d].reshape(d, d)
    if kind == 3:
        if is_diag(M):
            return diag_variances(1.0 / np.diag(M))
        out[:] = _chol_upper(M).reshape(-1)
        return NK_FULL, out
    if kind == 4:
        if is_diag(M):
            return diag_sigmas(1.0 / np.diag(M))
        out[:] = M.reshape(-1)
        return NK_FULL, out
    raise InvalidNoiseModel(f"unknown noise kind {kind}")


class Problem:
    """Container for the arrays above (missing buckets/types are simply absent)."""

    def __init__(self, arrays=None):
        self.a = dict(arrays or {})

    # ---- construction helpers -------------------------------------------------------------
    def set_variables(self, vtype, keys, state):
        name = VTYPES[vtype] if isinstance(vtype, int) else vtype
        t = VTYPES.index(name)
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        state = np.ascontiguousarray(state, dtype=np.float64).reshape(len(keys), VTYPE_STATE[t])
        self.a[name + "_keys"] = keys
        self.a[name] = state

    def set_factors(self, bucket, keys, z, nk, noise):
        name = BUCKETS[bucket] if isinstance(bucket, int) else bucket
        b = BUCKETS.index(name)
        m = len(z)
        d = BUCKET_DIM[b]
        keys = np.ascontiguousarray(keys, dtype=np.uint64).reshape(m, BUCKET_NKEYS[b])
        nk = np.broadcast_to(np.asarray(nk, dtype=np.int32), (m,)).copy()
        nz = np.zeros((m, d * d))
        noise = np.asarray(noise, dtype=np.float64)
        if noise.ndim == 1:
            noise = np.broadcast_to(noise, (m, noise.shape[0]))
        nz[:, : noise.shape[1]] = noise
        self.a[name + "_keys"] = keys
        self.a[name + "_z"] = np.ascontiguousarray(z, dtype=np.float64).reshape(m, BUCKET_Z[b])
        self.a[name + "_nk"] = nk
        self.a[name + "_noise"] = nz

    def set_robust(self, bucket, loss, k):
        """noiseModel::Robust::Create(mEstimator::<loss>::Create(k), base) for every factor of the bucket.
        Stored as `<bucket>_robust` = [loss id, k] (ids: LOSSES)."""
        name = BUCKETS[bucket] if isinstance(bucket, int) else bucket
        lid = LOSSES.index(loss) if isinstance(loss, str) else int(loss)
        self.a[name + "_robust"] = np.array([float(lid), float(k)])

    # ---- queries --------------------------------------------------------------------------
    def n_vars(self, t):
        k = self.a.get(VTYPES[t] + "_keys")
        return 0 if k is None else len(k)

    def n_factors(self, b):
        k = self.a.get(BUCKETS[b] + "_keys")
        return 0 if k is None else len(k)

    def graph_order(self):
        if "graph_order" in self.a:
            return np.asarray(self.a["graph_order"], dtype=np.int64)
        rows = [np.stack([np.full(self.n_factors(b), b, dtype=np.int64),
                          np.arange(self.n_factors(b), dtype=np.int64)], axis=1)
                for b in range(len(BUCKETS)) if self.n_factors(b)]
        return np.concatenate(rows, axis=0) if rows else np.zeros((0, 2), dtype=np.int64)

    def variables(self):
        """Global variable table sorted by key (the order of gtsam::Values, a std::map):
        keys[V], vtype[V], row[V] (row in its type array), dim[V]."""
        keys, vt, rows = [], [], []
        for t, name in enumerate(VTYPES):
            n = self.n_vars(t)
            if n:
                keys.append(self.a[name + "_keys"])
                vt.append(np.full(n, t, dtype=np.int32))
                rows.append(np.arange(n, dtype=np.int64))
        keys = np.concatenate(keys)
        vt = np.concatenate(vt)
        rows = np.concatenate(rows)
        o = np.argsort(keys, kind="stable")
        keys, vt, rows = keys[o], vt[o], rows[o]
        if len(keys) > 1 and np.any(keys[1:] == keys[:-1]):
            raise ValueError("duplicate keys in Values")
        dims = np.asarray(VTYPE_DIM, dtype=np.int32)[vt]
        return keys, vt, rows, dims

    def save(self, path):
        from . import gtbio
        gtbio.save(path, self.a)

    @staticmethod
    def load(path):
        from . import gtbio
        return Problem(gtbio.load(path))
