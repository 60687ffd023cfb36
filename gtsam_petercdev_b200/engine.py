"""Host-side handle on the CUDA engine: a thin, typed wrapper over the C ABI (include/gtsam_b200.h).

Method names follow the reference calls they replace:
  linearize()            NonlinearFactorGraph::linearize           nonlinear/NonlinearFactorGraph.cpp:239
  factorize(lambda)      buildDampedSystem + eliminateMultifrontal linear/GaussianFactorGraph.cpp:316
  backsolve()            GaussianBayesTree::optimize               linear/GaussianBayesTree.cpp:67
  linear_errors()        GaussianFactorGraph::error                linear/GaussianFactorGraph.cpp:71
  retract() / error()    Values::retract / NonlinearFactorGraph::error
  lm_optimize()          LevenbergMarquardtOptimizer::optimize
Every numeric method runs on the GPU; none has a host implementation.
"""
import ctypes as C
import numpy as np

from . import capi
from .problem import BUCKETS, BUCKET_DIM, VTYPES, Problem

_M = [3, 6, 2, 3, 6, 9, 3]
_W = [7, 13, 13, 4, 7, 10, 4]
_STATE_DEV = [4, 12, 17, 3]


class EngineError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"gtb status {status}: {msg}")
        self.status = status


class IndeterminantLinearSystemException(EngineError):
    """linear/linearExceptions.h:94 -- thrown by factorize() with the nearby variable's key."""

    def __init__(self, status, msg, key):
        super().__init__(status, msg)
        self.key = key


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Engine:
    def __init__(self, problem: Problem, device=0):
        self.lib = capi.lib()
        h = C.c_void_p()
        st = self.lib.gtb_create(int(device), C.byref(h))
        if st != capi.OK:
            raise EngineError(st, "gtb_create failed (no CUDA device?)" if device >= 0 else "gtb_create failed")
        self.h = h
        self.problem = problem
        self._keep = []
        a = problem.a
        for t, name in enumerate(VTYPES):
            if name + "_keys" in a:
                keys = np.ascontiguousarray(a[name + "_keys"], dtype=np.uint64)
                state = np.ascontiguousarray(a[name], dtype=np.float64)
                self._ck(self.lib.gtb_set_variables(self.h, t, len(keys), _ptr(keys), _ptr(state)))
        for b, name in enumerate(BUCKETS):
            if name + "_keys" in a:
                keys = np.ascontiguousarray(a[name + "_keys"], dtype=np.uint64)
                z = np.ascontiguousarray(a[name + "_z"], dtype=np.float64)
                nk = np.ascontiguousarray(a[name + "_nk"], dtype=np.int32)
                nz = np.ascontiguousarray(a[name + "_noise"], dtype=np.float64)
                assert nz.shape[1] == BUCKET_DIM[b] ** 2
                self._ck(self.lib.gtb_set_factors(self.h, b, len(z), _ptr(keys), _ptr(z), _ptr(nk), _ptr(nz)))
                if name + "_robust" in a:
                    self.set_robust(b, int(a[name + "_robust"][0]), float(a[name + "_robust"][1]))
        if "graph_order" in a:
            go = np.ascontiguousarray(a["graph_order"], dtype=np.int64)
            self._ck(self.lib.gtb_set_graph_order(self.h, len(go), _ptr(go)))
        self._ck(self.lib.gtb_finalize_problem(self.h))
        self.var_keys, self.var_type, self.var_row, self.var_dim = problem.variables()
        self.n_scalars = int(self.var_dim.sum())

    # ---- plumbing ---------------------------------------------------------------------------------
    def _ck(self, st):
        if st != capi.OK:
            raise EngineError(st, self.lib.gtb_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.gtb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- problem / plan ---------------------------------------------------------------------------
    def set_state(self, vtype, state):
        """Per-step host -> device refresh of one variable type (same layout as Problem arrays).  A page-locked array (e.g. a
        pinned torch tensor's .numpy()) is copied asynchronously from where it lies: keep it alive and unchanged until the next
        call that synchronises (trial(), error(), variables() ...)."""
        state = np.ascontiguousarray(state, dtype=np.float64)
        self._ck(self.lib.gtb_set_variables(self.h, vtype, len(state), None, _ptr(state)))

    def set_robust(self, bucket, loss, k):
        """noiseModel::Robust around the bucket's base noise (linear/NoiseModel.cpp:741); loss = GTB_LOSS_* id or name."""
        from .problem import LOSSES
        b = BUCKETS.index(bucket) if isinstance(bucket, str) else bucket
        lid = LOSSES.index(loss) if isinstance(loss, str) else int(loss)
        self._ck(self.lib.gtb_set_robust(self.h, b, lid, float(k)))

    def set_ordering(self, keys):
        if isinstance(keys, str):
            which = {"natural": 0, "schur": 1, "nested_dissection": 2}[keys]
            self._ck(self.lib.gtb_set_ordering_builtin(self.h, which))
            return
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        self._ck(self.lib.gtb_set_ordering(self.h, len(keys), _ptr(keys)))

    def set_distributed(self, rank, world, comm=None):
        """One engine per process / GPU.  `comm` is a capi.COMM_FN (see distributed.py); keep it alive with the engine."""
        self._comm = comm
        fn = C.cast(comm, C.c_void_p) if comm is not None else None
        self._ck(self.lib.gtb_set_distributed(self.h, int(rank), int(world), fn, None))

    def partition(self):
        nc = C.c_int64()
        self._ck(self.lib.gtb_plan_sizes(self.h, C.byref(nc), None, None, None))
        owner = np.zeros(nc.value, dtype=np.int32)
        st = np.zeros(16)
        self._ck(self.lib.gtb_partition_get(self.h, _ptr(owner), _ptr(st)))
        return dict(owner=owner, n_top=int(st[0]), reduced_doubles=int(st[1]), top_share=float(st[2]),
                    rank_share=st[3:].copy())

    def schedule_check(self):
        """Host-only invariants of the launch schedule (gtb_schedule_check); raises on a violation."""
        out = np.zeros(8, dtype=np.int64)
        self._ck(self.lib.gtb_schedule_check(self.h, _ptr(out)))
        names = ["launch_classes", "panel_step_launches", "schur_step_launches", "backsolve_launches", "inverse_blocks",
                 "pair_segments", "diag_segments", "step_tasks"]
        return dict(zip(names, out.tolist()))

    def plan(self):
        nc, nk, ncond, ns = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(self.lib.gtb_plan_sizes(self.h, C.byref(nc), C.byref(nk), C.byref(ncond), C.byref(ns)))
        nfrontal = np.zeros(nc.value, dtype=np.int64)
        nkeys = np.zeros(nc.value, dtype=np.int64)
        keys = np.zeros(nk.value, dtype=np.uint64)
        parent = np.zeros(nc.value, dtype=np.int64)
        self._ck(self.lib.gtb_plan_get(self.h, _ptr(nfrontal), _ptr(nkeys), _ptr(keys), _ptr(parent)))
        stats = np.zeros(16)
        self._ck(self.lib.gtb_plan_stats(self.h, _ptr(stats)))
        names = ["levels", "max_front", "flops", "arena_doubles", "n_small", "n_large", "flops_large", "assembly_bytes",
                 "flops_syrk_frontal", "flops_syrk_schur", "bytes_backsolve", "bytes_small", "bytes_linearize", "bytes_thin", "bytes_point", "bytes_dense_small"]
        return dict(clique_nfrontal=nfrontal, clique_nkeys=nkeys, clique_keys=keys, clique_parent=parent,
                    n_conditional_doubles=ncond.value, n_scalars=ns.value, stats=dict(zip(names, stats.tolist())))

    # ---- stages -----------------------------------------------------------------------------------
    def linearize(self):
        self._ck(self.lib.gtb_linearize(self.h))

    def jacobians(self, bucket):
        b = BUCKETS.index(bucket) if isinstance(bucket, str) else bucket
        n = self.problem.n_factors(b)
        out = np.zeros((n, _M[b] * _W[b]))
        self._ck(self.lib.gtb_get_jacobians(self.h, b, _ptr(out)))
        return out

    def hessian_diagonal(self):
        out = np.zeros(self.n_scalars)
        self._ck(self.lib.gtb_hessian_diagonal(self.h, _ptr(out)))
        return out

    def factorize(self, lam=0.0, diagonal_damping=False, min_diagonal=1e-6, max_diagonal=1e32):
        key = C.c_uint64(0)
        st = self.lib.gtb_factorize(self.h, float(lam), int(bool(diagonal_damping)), float(min_diagonal),
                                    float(max_diagonal), C.byref(key))
        if st == capi.ERR_INDETERMINATE:
            raise IndeterminantLinearSystemException(st, self.lib.gtb_last_error(self.h).decode(), key.value)
        self._ck(st)

    def trial(self, lam=0.0, diagonal_damping=False, min_diagonal=1e-6, max_diagonal=1e32):
        """One tryLambda worth of device work with one host sync: (linear error at 0, at delta, error of the trial values)."""
        out = np.zeros(3)
        key = C.c_uint64(0)
        st = self.lib.gtb_trial(self.h, float(lam), int(bool(diagonal_damping)), float(min_diagonal), float(max_diagonal),
                                _ptr(out), C.byref(key))
        if st == capi.ERR_INDETERMINATE:
            raise IndeterminantLinearSystemException(st, self.lib.gtb_last_error(self.h).decode(), key.value)
        self._ck(st)
        return float(out[0]), float(out[1]), float(out[2])

    def backsolve(self):
        self._ck(self.lib.gtb_backsolve(self.h))

    def delta(self):
        out = np.zeros(self.n_scalars)
        self._ck(self.lib.gtb_get_delta(self.h, _ptr(out)))
        return out

    def conditionals(self):
        n = self.plan()["n_conditional_doubles"]
        out = np.zeros(n)
        self._ck(self.lib.gtb_get_conditionals(self.h, _ptr(out)))
        return out

    def linear_errors(self):
        e0, e1 = C.c_double(), C.c_double()
        self._ck(self.lib.gtb_linear_errors(self.h, C.byref(e0), C.byref(e1)))
        return e0.value, e1.value

    def retract(self):
        self._ck(self.lib.gtb_retract(self.h))

    def error(self, which=0):
        e = C.c_double()
        self._ck(self.lib.gtb_error(self.h, int(which), C.byref(e)))
        return e.value

    def accept(self):
        self._ck(self.lib.gtb_accept(self.h))

    def variables(self, vtype, which=0):
        t = VTYPES.index(vtype) if isinstance(vtype, str) else vtype
        out = np.zeros((self.problem.n_vars(t), _STATE_DEV[t]))
        self._ck(self.lib.gtb_get_variables(self.h, t, int(which), _ptr(out)))
        return out

    # ---- optimizers -------------------------------------------------------------------------------
    def lm_params(self, ceres=False, **kw):
        p = capi.LMParams()
        (self.lib.gtb_lm_params_ceres if ceres else self.lib.gtb_lm_params_default)(C.byref(p))
        for k, v in kw.items():
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
        return p

    def lm_optimize(self, params=None, trace_rows=4096):
        p = params or self.lm_params()
        r = capi.LMResult()
        trace = np.zeros((trace_rows, 5))
        self._ck(self.lib.gtb_lm_optimize(self.h, C.byref(p), C.byref(r), _ptr(trace), trace_rows))
        return dict(initial_error=r.initial_error, final_error=r.final_error, final_lambda=r.final_lambda,
                    iterations=r.iterations, inner_iterations=r.inner_iterations, trace=trace[: r.log_rows].copy())

    def lm_iterate(self, params, lam, factor):
        l, f, e, inner = C.c_double(lam), C.c_double(factor), C.c_double(), C.c_int32()
        self._ck(self.lib.gtb_lm_iterate(self.h, C.byref(params), C.byref(l), C.byref(f), C.byref(e), C.byref(inner)))
        return l.value, f.value, e.value, inner.value

    def gn_optimize(self, max_iterations=100, relative_error_tol=1e-5, absolute_error_tol=1e-5):
        r = capi.LMResult()
        self._ck(self.lib.gtb_gn_optimize(self.h, int(max_iterations), float(relative_error_tol),
                                          float(absolute_error_tol), C.byref(r)))
        return dict(initial_error=r.initial_error, final_error=r.final_error, iterations=r.iterations)

    # ---- instrumentation --------------------------------------------------------------------------
    def marginal_covariances(self, keys):
        """Marginals(graph, values).marginalCovariance(key) for every key (nonlinear/Marginals.cpp:108): list of d x d
        arrays.  Needs linearize() + factorize(0.0) on the current values."""
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        self._ck(self.lib.gtb_marginals_compute(self.h))
        _, vt, _, dims = self.problem.variables()
        kk, _, _, _ = self.problem.variables()
        dim_of = dict(zip(kk.tolist(), dims.tolist()))
        ds = [dim_of.get(int(k), 0) for k in keys]
        out = np.zeros(int(sum(d * d for d in ds)) + 1)
        self._ck(self.lib.gtb_marginal_covariance(self.h, len(keys), _ptr(keys), _ptr(out)))
        res, o = [], 0
        for d in ds:
            res.append(out[o:o + d * d].reshape(d, d).copy())
            o += d * d
        return res

    def kernel_launches(self):
        return int(self.lib.gtb_kernel_launches(self.h))

    def stage_ms(self):
        out = np.zeros(8)
        self._ck(self.lib.gtb_stage_ms(self.h, _ptr(out)))
        names = ["linearize", "factorize", "backsolve", "linear_errors", "retract", "error", "large_fronts", "hessian_diagonal"]
        return dict(zip(names, out.tolist()))

    def set_stage_timing(self, on):
        self._ck(self.lib.gtb_set_stage_timing(self.h, int(bool(on))))

    def profile(self, on=True):
        self._ck(self.lib.gtb_profile(self.h, int(bool(on))))

    def profile_read(self):
        """{kernel class: (total ms, launches)} since the last read, from per-launch CUDA events on the engine stream."""
        n = self.lib.gtb_profile_classes()
        ms, cnt = np.zeros(n), np.zeros(n, dtype=np.int64)
        self._ck(self.lib.gtb_profile_read(self.h, _ptr(ms), _ptr(cnt)))
        return {self.lib.gtb_profile_class_name(k).decode(): (float(ms[k]), int(cnt[k])) for k in range(n) if cnt[k]}

    def profile_read_raw(self, cap=1 << 20):
        """(class names per launch, ms per launch) in launch order since the last read."""
        k = np.zeros(cap, dtype=np.int32)
        ms = np.zeros(cap, dtype=np.float32)
        n = self.lib.gtb_profile_read_raw(self.h, _ptr(k), _ptr(ms), cap)
        names = [self.lib.gtb_profile_class_name(i).decode() for i in range(self.lib.gtb_profile_classes())]
        return [names[i] for i in k[:n]], ms[:n].copy()

    def stream(self):
        return self.lib.gtb_stream(self.h)
