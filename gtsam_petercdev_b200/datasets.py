"""Synthetic workloads of the shapes BASELINE.json names and readers for the
reference's dataset formats (g2o / TORO EDGE2/EDGE3, BAL).  Everything is seeded and vectorised.

C1  m3500_shape()   3,500 Pose2, 3,499 odometry + 1,951 loop closures, prior on pose 0
C2  sphere_shape()  50 x 50 Pose3 on a sphere, ring + meridian edges, prior on pose 0
C3  ba_shape(1723, 156000)   Ladybug-scale bundle adjustment (~679k GeneralSFMFactor<Cal3Bundler>)
C4  ba_shape(1778, 993000, mean_obs=5.03)  Venice-scale
"""
import os

import numpy as np
from .problem import Problem, symbol


# --------------------------------------------------------------------------------------------
# small geometry helpers used only to fabricate data (not the product arithmetic)
# --------------------------------------------------------------------------------------------
def _rodrigues(w):
    """Rotation matrices exp([w]x) for w[n,3]."""
    w = np.asarray(w, dtype=np.float64)
    th = np.linalg.norm(w, axis=1)
    small = th < 1e-12
    ths = np.where(small, 1.0, th)
    k = w / ths[:, None]
    K = np.zeros((len(w), 3, 3))
    K[:, 0, 1], K[:, 0, 2] = -k[:, 2], k[:, 1]
    K[:, 1, 0], K[:, 1, 2] = k[:, 2], -k[:, 0]
    K[:, 2, 0], K[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = np.sin(th)[:, None, None], np.cos(th)[:, None, None]
    R = np.eye(3)[None] + s * K + (1 - c) * (K @ K)
    R[small] = np.eye(3)
    return R


def _wrap(a):
    return np.arctan2(np.sin(a), np.cos(a))


def _pose2_between(a, b):
    """a^-1 * b for (x, y, theta) rows."""
    c, s = np.cos(a[:, 2]), np.sin(a[:, 2])
    dx, dy = b[:, 0] - a[:, 0], b[:, 1] - a[:, 1]
    return np.stack([c * dx + s * dy, -s * dx + c * dy, _wrap(b[:, 2] - a[:, 2])], axis=1)


def _pose2_compose(a, d):
    c, s = np.cos(a[:, 2]), np.sin(a[:, 2])
    return np.stack([a[:, 0] + c * d[:, 0] - s * d[:, 1], a[:, 1] + s * d[:, 0] + c * d[:, 1],
                     _wrap(a[:, 2] + d[:, 2])], axis=1)


# --------------------------------------------------------------------------------------------
# C1: Manhattan-world Pose2 graph
# --------------------------------------------------------------------------------------------
def m3500_shape(n_poses=3500, n_closures=1951, seed=3500):
    rng = np.random.default_rng(seed)
    turn = rng.choice([0.0, np.pi / 2, -np.pi / 2], size=n_poses - 1, p=[0.7, 0.15, 0.15])
    theta = np.concatenate([[0.0], np.cumsum(turn)])
    # turn, then step 1 m along the new heading
    xy = np.zeros((n_poses, 2))
    xy[1:, 0] = np.cumsum(np.cos(theta[1:]))
    xy[1:, 1] = np.cumsum(np.sin(theta[1:]))
    truth = np.stack([xy[:, 0], xy[:, 1], _wrap(theta)], axis=1)

    # loop closures: pairs closer than 3 m and more than 20 steps apart
    grid = {}
    cell = np.floor(xy / 3.0).astype(np.int64)
    for i, (cx, cy) in enumerate(cell):
        grid.setdefault((cx, cy), []).append(i)
    cand = []
    for i, (cx, cy) in enumerate(cell):
        for ox in (-1, 0, 1):
            for oy in (-1, 0, 1):
                for j in grid.get((cx + ox, cy + oy), ()):
                    if j > i + 20 and np.hypot(*(xy[i] - xy[j])) < 3.0:
                        cand.append((i, j))
    cand = np.asarray(cand, dtype=np.int64).reshape(-1, 2)
    if len(cand) < n_closures:
        raise RuntimeError(f"only {len(cand)} closure candidates")
    cand = cand[rng.choice(len(cand), size=n_closures, replace=False)]
    cand = cand[np.lexsort((cand[:, 1], cand[:, 0]))]

    odo = np.stack([np.arange(n_poses - 1), np.arange(1, n_poses)], axis=1)
    ij = np.concatenate([odo, cand], axis=0)
    z = _pose2_between(truth[ij[:, 0]], truth[ij[:, 1]])
    # SURVEY 8d suggests sigma (0.15, 0.15, 0.1); with that much odometry drift the reference's own LM
    # rejects every lambda at iteration 0, so the shipped shape uses a convergent noise level.
    sig = np.array([0.05, 0.05, 0.02])
    z = z + rng.normal(size=z.shape) * sig
    z[:, 2] = _wrap(z[:, 2])

    m = len(ij)
    info = np.zeros((m, 3, 3))
    info[:] = np.diag(1.0 / sig ** 2)
    # closures carry a full (rotated) information matrix so the Gaussian-R whitening path is exercised
    a = rng.uniform(0, np.pi, size=n_closures)
    ca, sa = np.cos(a), np.sin(a)
    Rot = np.zeros((n_closures, 3, 3))
    Rot[:, 0, 0], Rot[:, 0, 1], Rot[:, 1, 0], Rot[:, 1, 1], Rot[:, 2, 2] = ca, -sa, sa, ca, 1.0
    D = np.diag([1 / 0.1 ** 2, 1 / 0.2 ** 2, 1 / 0.1 ** 2])
    info[n_poses - 1:] = Rot @ D @ Rot.transpose(0, 2, 1)

    # initial values: chained noisy odometry
    init = np.zeros((n_poses, 3))
    zo = z[: n_poses - 1]
    for i in range(n_poses - 1):
        init[i + 1] = _pose2_compose(init[i:i + 1], zo[i:i + 1])[0]

    p = Problem()
    keys = np.arange(n_poses, dtype=np.uint64)
    p.set_variables("pose2", keys, init)
    p.set_factors("between2", keys[ij], z, 3, info.reshape(m, 9))
    p.set_factors("prior2", keys[:1], init[:1], 5, np.array([[1e-6, 1e-6, 1e-8]]))
    p.truth = truth
    return p


# --------------------------------------------------------------------------------------------
# C2: Pose3 sphere
# --------------------------------------------------------------------------------------------
def _pose3_rows(R, t):
    return np.concatenate([R.reshape(len(R), 9), t], axis=1)


def _pose3_between(Ra, ta, Rb, tb):
    Rt = Ra.transpose(0, 2, 1)
    return Rt @ Rb, np.einsum("nij,nj->ni", Rt, tb - ta)


def sphere_shape(rings=50, per_ring=50, radius=50.0, sigma_R=0.05, sigma_t=0.1, seed=2500):
    rng = np.random.default_rng(seed)
    n = rings * per_ring
    ri, pi_ = np.divmod(np.arange(n), per_ring)
    lat = (ri + 1) / (rings + 1) * np.pi - np.pi / 2
    lon = pi_ / per_ring * 2 * np.pi
    t = radius * np.stack([np.cos(lat) * np.cos(lon), np.cos(lat) * np.sin(lon), np.sin(lat)], axis=1)
    # orientation: x tangent along the ring, z outward
    zax = t / radius
    xax = np.stack([-np.sin(lon), np.cos(lon), np.zeros(n)], axis=1)
    yax = np.cross(zax, xax)
    R = np.stack([xax, yax, zax], axis=2)

    ring_edges = np.stack([np.arange(n), ri * per_ring + (pi_ + 1) % per_ring], axis=1)
    mer = np.arange(n - per_ring)
    mer_edges = np.stack([mer, mer + per_ring], axis=1)
    # graph order: odometry-like chain first (i -> i+1 over the flattened index), then the rest
    chain = np.stack([np.arange(n - 1), np.arange(1, n)], axis=1)
    extra = np.concatenate([ring_edges[pi_ == per_ring - 1], mer_edges], axis=0)
    ij = np.concatenate([chain, extra], axis=0)

    Rz, tz = _pose3_between(R[ij[:, 0]], t[ij[:, 0]], R[ij[:, 1]], t[ij[:, 1]])
    m = len(ij)
    Rz = Rz @ _rodrigues(rng.normal(size=(m, 3)) * sigma_R)
    tz = tz + rng.normal(size=(m, 3)) * sigma_t
    # g2o EDGE3-like full information (slam/dataset.cpp:819-859), mildly correlated
    base = np.diag(np.concatenate([np.full(3, 1 / sigma_R ** 2), np.full(3, 1 / sigma_t ** 2)]))
    L = np.eye(6) + 0.05 * np.triu(np.ones((6, 6)), 1)
    info = np.broadcast_to(L.T @ base @ L, (m, 6, 6))

    # initial values: chained odometry from the first pose
    Ri, ti = np.zeros((n, 3, 3)), np.zeros((n, 3))
    Ri[0], ti[0] = R[0], t[0]
    for i in range(n - 1):
        ti[i + 1] = ti[i] + Ri[i] @ tz[i]
        Ri[i + 1] = Ri[i] @ Rz[i]

    p = Problem()
    keys = np.arange(n, dtype=np.uint64)
    p.set_variables("pose3", keys, _pose3_rows(Ri, ti))
    p.set_factors("between3", keys[ij], _pose3_rows(Rz, tz), 3, info.reshape(m, 36))
    p.set_factors("prior3", keys[:1], _pose3_rows(Ri[:1], ti[:1]), 5,
                  np.array([[1e-6, 1e-6, 1e-6, 1e-4, 1e-4, 1e-4]]))
    return p


# --------------------------------------------------------------------------------------------
# C3 / C4: bundle adjustment on a closed street-like loop
# --------------------------------------------------------------------------------------------
def _project(Rc, tc, cal, pts):
    """PinholeCamera<Cal3Bundler>::project for matched rows (no Jacobians); returns (uv, depth)."""
    q = np.einsum("nji,nj->ni", Rc, pts - tc)          # R^T (p - t)
    x, y = q[:, 0] / q[:, 2], q[:, 1] / q[:, 2]
    r = x * x + y * y
    g = 1.0 + (cal[:, 1] + cal[:, 2] * r) * r
    f = cal[:, 0]
    return np.stack([cal[:, 3] + f * g * x, cal[:, 4] + f * g * y], axis=1), q[:, 2]


def ba_shape(n_cams=1723, n_points=156000, mean_obs=4.35, window=60, long_range=0.0,
             spacing=1.0, seed=1723, pixel_sigma=0.5, cam_sigma=0.005, point_sigma=0.05):
    rng = np.random.default_rng(seed)
    radius = n_cams * spacing / (2 * np.pi)
    ang = np.arange(n_cams) / n_cams * 2 * np.pi
    fwd = np.stack([np.cos(ang), np.sin(ang), np.zeros(n_cams)], axis=1)      # camera z: outward
    down = np.broadcast_to(np.array([0.0, 0.0, -1.0]), (n_cams, 3))            # camera y
    right = np.cross(down, fwd)                                                # camera x
    Rc = np.stack([right, down, fwd], axis=2)
    tc = radius * fwd
    cal = np.stack([rng.uniform(600, 1000, n_cams), rng.normal(0, 1e-7, n_cams),
                    rng.normal(0, 1e-13, n_cams), np.zeros(n_cams), np.zeros(n_cams)], axis=1)

    half = window // 2
    home = np.sort(rng.integers(0, n_cams, size=n_points))
    depth = rng.uniform(10.0, 40.0, n_points)
    lat = rng.uniform(-3.0, 3.0, n_points)
    hgt = rng.uniform(-3.0, 3.0, n_points)
    pts = tc[home] + depth[:, None] * fwd[home] + lat[:, None] * right[home] - hgt[:, None] * down[home]

    p_geo = 1.0 / (1.0 + (mean_obs - 2.0))
    k = np.minimum(2 + rng.geometric(p_geo, size=n_points) - 1, 2 * half + 1)
    cam_idx, pt_idx = [], []
    chunk = 200000
    for s in range(0, n_points, chunk):
        e = min(s + chunk, n_points)
        score = rng.random((e - s, 2 * half + 1))
        order = np.argsort(score, axis=1)
        mask = np.arange(2 * half + 1)[None, :] < k[s:e, None]
        offs = order - half
        rows = np.broadcast_to(np.arange(s, e)[:, None], offs.shape)
        cam_idx.append(((home[s:e, None] + offs) % n_cams)[mask])
        pt_idx.append(rows[mask])
    cam_idx = np.concatenate(cam_idx)
    pt_idx = np.concatenate(pt_idx)
    if long_range > 0:
        sel = np.nonzero(rng.random(n_points) < long_range)[0]
        cam_idx = np.concatenate([cam_idx, rng.integers(0, n_cams, size=len(sel))])
        pt_idx = np.concatenate([pt_idx, sel])
    o = np.lexsort((cam_idx, pt_idx))
    cam_idx, pt_idx = cam_idx[o], pt_idx[o]
    keep = np.ones(len(o), dtype=bool)
    keep[1:] = (cam_idx[1:] != cam_idx[:-1]) | (pt_idx[1:] != pt_idx[:-1])
    cam_idx, pt_idx = cam_idx[keep], pt_idx[keep]

    uv, z = _project(Rc[cam_idx], tc[cam_idx], cal[cam_idx], pts[pt_idx])
    ok = z > 0.5
    cam_idx, pt_idx, uv = cam_idx[ok], pt_idx[ok], uv[ok]
    # every variable must appear in the graph (the reference's COLAMD ordering is built from the graph, and LM damping
    # adds a prior per Values entry): drop points left with fewer than two observations and renumber
    cnt = np.bincount(pt_idx, minlength=n_points)
    keep_pt = cnt >= 2
    sel = keep_pt[pt_idx]
    cam_idx, pt_idx, uv = cam_idx[sel], pt_idx[sel], uv[sel]
    remap = np.cumsum(keep_pt) - 1
    pt_idx = remap[pt_idx]
    pts = pts[keep_pt]
    n_points = len(pts)
    uv = uv + rng.normal(size=uv.shape) * pixel_sigma

    # initial estimate: truth perturbed on the pose tangent and on the points
    dR = _rodrigues(rng.normal(size=(n_cams, 3)) * cam_sigma)
    Ri = Rc @ dR
    ti = tc + np.einsum("nij,nj->ni", Rc, rng.normal(size=(n_cams, 3)) * cam_sigma)
    pi_ = pts + rng.normal(size=pts.shape) * point_sigma
    cams = np.concatenate([Ri.reshape(n_cams, 9), ti, cal], axis=1)

    ckeys = symbol("c", np.arange(n_cams))
    pkeys = symbol("p", np.arange(n_points))
    p = Problem()
    p.set_variables("cam", ckeys, cams)
    p.set_variables("point", pkeys, pi_)
    # Isotropic::Sigma(2, 1.0) -> Unit under the smart rule (examples/SFMExample_bal.cpp:52)
    p.set_factors("sfm", np.stack([ckeys[cam_idx], pkeys[pt_idx]], axis=1), uv, 1, np.array([[1.0]]))
    p.set_factors("priorcam", ckeys[:1], cams[:1], 1, np.array([[0.1]]))
    p.set_factors("priorpt", pkeys[:1], pi_[:1], 1, np.array([[0.1]]))
    return p


def ladybug_shape(seed=1723):
    return ba_shape(1723, 156000, 4.35, seed=seed)


def venice_shape(seed=1778):
    return ba_shape(1778, 993000, 5.03, seed=seed)


def read_bal(path):
    """Bundle Adjustment in the Large file -> Problem, following SfmData::FromBalFile (sfm/SfmData.cpp:189-247) and the
    graph of tests/testGeneralSFMFactorB.cpp:44-58: every number is read as a 32-bit float, v is negated, the OpenGL
    camera (R, t) becomes wRc = R^T diag(1,-1,-1), wtc = -R^T t (openGL2gtsam, sfm/SfmData.cpp:79-86), cameras get the
    plain keys 0..n-1 and points the keys P(j), measurements get Unit noise, no priors."""
    tok = open(path).read().split()
    n_cams, n_points, n_obs = int(tok[0]), int(tok[1]), int(tok[2])
    f32 = lambda a: np.array(a, dtype=np.float32).astype(np.float64)   # noqa: E731
    pos = 3
    obs = np.array(tok[pos:pos + 4 * n_obs]).reshape(n_obs, 4)
    pos += 4 * n_obs
    cam_idx, pt_idx = obs[:, 0].astype(np.int64), obs[:, 1].astype(np.int64)
    uv = np.stack([f32(obs[:, 2]), -f32(obs[:, 3])], axis=1)
    cam = f32(tok[pos:pos + 9 * n_cams]).reshape(n_cams, 9)
    pos += 9 * n_cams
    pts = f32(tok[pos:pos + 3 * n_points]).reshape(n_points, 3)
    R = _rodrigues(cam[:, 0:3])
    wRc = np.transpose(R, (0, 2, 1)) @ np.diag([1.0, -1.0, -1.0])
    wtc = -np.einsum("nji,nj->ni", R, cam[:, 3:6])
    cal = np.concatenate([cam[:, 6:9], np.zeros((n_cams, 2))], axis=1)
    cams = np.concatenate([wRc.reshape(n_cams, 9), wtc, cal], axis=1)
    # the reference iterates tracks (points) and, inside, their measurements in file order
    o = np.argsort(pt_idx, kind="stable")
    ckeys = np.arange(n_cams, dtype=np.uint64)
    pkeys = symbol("p", np.arange(n_points))
    p = Problem()
    p.set_variables("cam", ckeys, cams)
    p.set_variables("point", pkeys, pts)
    p.set_factors("sfm", np.stack([ckeys[cam_idx[o]], pkeys[pt_idx[o]]], axis=1), uv[o], 1, np.array([[1.0]]))
    return p


# ---- native ingestion (csrc/ingest.cpp behind the C ABI): files straight to the flat tables -----------------------
class DatasetError(ValueError):
    pass


def _dataset_to_problem(h):
    import ctypes as C
    from . import capi
    from .problem import BUCKETS, BUCKET_Z, BUCKET_DIM, BUCKET_NKEYS, VTYPES, VTYPE_STATE, LOSSES
    lib = capi.lib()

    def arr(ptr, n, ctype, dtype):
        if n == 0 or not ptr:
            return np.zeros(0, dtype=dtype)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(n,)).astype(dtype, copy=True)
    p = Problem()
    for t, name in enumerate(VTYPES):
        n = lib.gtb_dataset_num_variables(h, t)
        if n > 0:
            p.set_variables(name, arr(lib.gtb_dataset_variable_keys(h, t), n, C.c_uint64, np.uint64),
                            arr(lib.gtb_dataset_variable_state(h, t), n * VTYPE_STATE[t], C.c_double, np.float64).reshape(n, VTYPE_STATE[t]))
    for b, name in enumerate(BUCKETS):
        n = lib.gtb_dataset_num_factors(h, b)
        if n > 0:
            d = BUCKET_DIM[b]
            p.set_factors(name, arr(lib.gtb_dataset_factor_keys(h, b), n * BUCKET_NKEYS[b], C.c_uint64, np.uint64).reshape(n, BUCKET_NKEYS[b]),
                          arr(lib.gtb_dataset_factor_z(h, b), n * BUCKET_Z[b], C.c_double, np.float64).reshape(n, BUCKET_Z[b]),
                          arr(lib.gtb_dataset_factor_noise_kind(h, b), n, C.c_int32, np.int32),
                          arr(lib.gtb_dataset_factor_noise(h, b), n * d * d, C.c_double, np.float64).reshape(n, d * d))
    k = C.c_double(0.0)
    loss = lib.gtb_dataset_robust(h, C.byref(k))
    if loss:
        p.set_robust("between2", LOSSES[loss], k.value)
    return p


def _read_native(fn, *args):
    import ctypes as C
    from . import capi
    lib = capi.lib()
    h = C.c_void_p()
    st = getattr(lib, fn)(*args, C.byref(h))
    if st != capi.OK:
        raise DatasetError(lib.gtb_dataset_last_error().decode())
    try:
        return _dataset_to_problem(h)
    finally:
        lib.gtb_dataset_free(h)


def read_g2o(path, is3d=False, kernel="none"):
    """gtsam::readG2o(file, is3D, kernelFunctionType) (slam/dataset.cpp:621-633) -> Problem, parsed by the native
    reader (gtb_dataset_read_g2o).  kernel: "none" | "huber" | "tukey"."""
    from .problem import LOSSES
    return _read_native("gtb_dataset_read_g2o", os.fsencode(path), int(bool(is3d)), LOSSES.index(kernel))


def sphere2500_file(path=None):
    """C2 on the dataset the reference ships (examples/Data/sphere2500.txt, a copy travels under tests/golden/data/).  The file
    holds TORO EDGE3 records only (no vertices): the factors come from the native reader (bit-exact against the reference's
    load3D, tests/test_ingest_sweep_cpu.py); the initial values are dead reckoning along the odometry edges (i, i + 1) from the
    identity, which is what the reference's own example for this file does (matlab/gtsam_examples/Pose3SLAMExample_graph.m:
    load3D(..., smart = true) chains the odometry); pose 0 is held by the prior of examples/Pose3SLAMExample_g2o.cpp:50-53,
    noiseModel::Diagonal::Variances(1e-6 x3, 1e-4 x3)."""
    if path is None:
        path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "data", "sphere2500.txt")
    p = read_g2o(path, is3d=True)
    fk, z = p.a["between3_keys"], p.a["between3_z"]
    keys = np.unique(fk)
    pose = {int(keys[0]): (np.eye(3), np.zeros(3))}
    for (a, b), zz in zip(fk, z):
        a, b = int(a), int(b)
        if b == a + 1 and a in pose and b not in pose:
            Ra, ta = pose[a]
            pose[b] = (Ra @ zz[:9].reshape(3, 3), ta + Ra @ zz[9:])
    if len(pose) != len(keys):
        raise DatasetError("sphere2500: the odometry chain does not reach every pose")
    rows = np.stack([np.concatenate([pose[int(k)][0].reshape(9), pose[int(k)][1]]) for k in keys])
    p.set_variables("pose3", keys, rows)
    p.set_factors("prior3", keys[:1], rows[:1], 5, np.array([[1e-6, 1e-6, 1e-6, 1e-4, 1e-4, 1e-4]]))   # 5 = GTB_NOISE_DIAG_VARIANCES
    return p


def read_bal_native(path):
    """SfmData::FromBalFile + the projection factors of SFMExample_bal (gtb_dataset_read_bal) -> Problem."""
    return _read_native("gtb_dataset_read_bal", os.fsencode(path))


def write_g2o(path, problem):
    """gtsam::writeG2o (slam/dataset.cpp:636-735) for the poses and between factors of a Problem."""
    from . import capi
    from .problem import resolve_noise
    lib = capi.lib()
    a = problem.a

    def get(name, dtype, cols):
        v = a.get(name)
        if v is None or len(v) == 0:
            return np.zeros((0, cols), dtype=dtype)
        return np.ascontiguousarray(v, dtype=dtype).reshape(len(v), cols)

    def info(bucket, d):
        nk, noise = a.get(bucket + "_nk"), a.get(bucket + "_noise")
        if nk is None or len(nk) == 0:
            return np.zeros((0, d * d))
        out = np.zeros((len(nk), d * d))
        for i in range(len(nk)):
            kind, R = resolve_noise(int(nk[i]), np.asarray(noise[i], dtype=np.float64), d)
            Rm = np.eye(d) if kind == 0 else (np.eye(d) * R[0] if kind == 1 else (np.diag(R[:d]) if kind == 2 else np.asarray(R).reshape(d, d)))
            out[i] = (Rm.T @ Rm).reshape(-1)
        return out
    p2k, p2 = get("pose2_keys", np.uint64, 1), get("pose2", np.float64, 3)
    p3k, p3 = get("pose3_keys", np.uint64, 1), get("pose3", np.float64, 12)
    b2k, b2z, b2i = get("between2_keys", np.uint64, 2), get("between2_z", np.float64, 3), info("between2", 3)
    b3k, b3z, b3i = get("between3_keys", np.uint64, 2), get("between3_z", np.float64, 12), info("between3", 6)
    ptr = lambda x: x.ctypes.data if x.size else None   # noqa: E731
    st = lib.gtb_write_g2o(os.fsencode(path), len(p2k), ptr(p2k), ptr(p2), len(p3k), ptr(p3k), ptr(p3),
                           len(b2k), ptr(b2k), ptr(b2z), ptr(b2i), len(b3k), ptr(b3k), ptr(b3z), ptr(b3i))
    if st != capi.OK:
        raise DatasetError(lib.gtb_dataset_last_error().decode())
