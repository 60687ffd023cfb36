"""gtb_bucket_error_noalloc (include/gtsam_b200.h): the allocation-free variant of the nonlinear-error pass, whose
scratch is a fixed file-scope __device__ array.  No counterpart in the reference;
checked here against the existing entry point gtb_error through the C ABI, on two calls in a row.

Collected last on purpose: a failure here must not hide any other test.  Passes alone in a
fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import golden_problem, load_golden
from gtsam_petercdev_b200 import capi
from gtsam_petercdev_b200.problem import (BUCKETS, BUCKET_DIM, BUCKET_NKEYS, BUCKET_VTYPES, NK_UNIT, VTYPES, resolve_noise)

REL_TOL = 1e-12   # the tolerance of the error comparisons in tests/test_parity_gpu.py (summation order differs only)


def _device_state(vtype, host):
    """Caller-side copy of the device layout: Pose2 (x, y, theta) -> (cos, sin, x, y); the rest unchanged."""
    host = np.asarray(host, dtype=np.float64)
    if vtype == 0:
        return np.stack([np.cos(host[:, 2]), np.sin(host[:, 2]), host[:, 0], host[:, 1]], axis=1)
    return host


class _Buckets:
    """Device copies (torch tensors: plumbing for device memory only) of everything the entry point reads."""

    def __init__(self, problem, torch, dev):
        a = problem.a
        self.torch, self.dev = torch, dev
        self.rows = {}
        self.items = []
        for b, name in enumerate(BUCKETS):
            if name + "_keys" not in a:
                continue
            keys = np.asarray(a[name + "_keys"], dtype=np.uint64).reshape(-1, BUCKET_NKEYS[b])
            rows = []
            for k, vt in enumerate(BUCKET_VTYPES[b]):
                vkeys = np.asarray(a[VTYPES[vt] + "_keys"], dtype=np.uint64)
                order = np.argsort(vkeys, kind="stable")
                pos = np.searchsorted(vkeys[order], keys[:, k])
                assert np.array_equal(vkeys[order][pos], keys[:, k])
                rows.append(self._up(order[pos].astype(np.int32)))
            z = np.asarray(a[name + "_z"], dtype=np.float64)
            if b in (0, 3):
                z = _device_state(0, z)
            d = BUCKET_DIM[b]
            res = [resolve_noise(int(k), nz, d) for k, nz in zip(a[name + "_nk"], a[name + "_noise"])]
            nk = np.array([r[0] for r in res], dtype=np.int8)
            noise = np.stack([r[1] for r in res]).astype(np.float64)
            unit = bool(np.all(nk == NK_UNIT))
            self.items.append((b, len(z), rows, self._up(z), None if unit else self._up(nk), None if unit else self._up(noise)))
        self.out = torch.zeros(len(self.items), dtype=torch.float64, device=dev)

    def _up(self, arr):
        return self.torch.from_numpy(np.ascontiguousarray(arr)).to(self.dev)

    def total(self, lib, states, stream):
        """Sum over buckets of the no-alloc entry point's result; all launches on `stream`, one read-back at the end."""
        for i, (b, n, rows, z, nk, noise) in enumerate(self.items):
            vts = BUCKET_VTYPES[b]
            p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
            st = lib.gtb_bucket_error_noalloc(0, b, n, p(states[vts[0]]), p(states[vts[1]]) if len(vts) > 1 else None,
                                              p(rows[0]), p(rows[1]) if len(rows) > 1 else None, p(z), p(nk), p(noise),
                                              C.c_void_p(self.out.data_ptr() + 8 * i), C.c_void_p(stream))
            assert st == capi.OK, lib.gtb_noalloc_last_error().decode()
        self.torch.cuda.synchronize()
        return float(self.out.cpu().numpy().sum())


@pytest.mark.gpu
@pytest.mark.parametrize("name,tag", [("ba_30_500", "ba_30_500__solve_metis_0.0001_1"),
                                      ("pose3_120", "pose3_120__solve_metis_0.001_1"),
                                      ("pose2_300", "pose2_300__solve_metis_0.001_0")])
def test_noalloc_error_matches_gtb_error_on_two_calls_in_a_row(name, tag):
    import torch
    from gtsam_petercdev_b200.engine import Engine
    lib = capi.lib()
    dev = torch.device("cuda:0")
    problem = golden_problem(name)
    g = load_golden(tag)
    e = Engine(problem, device=0)
    try:
        e.set_ordering(g["ordering"])
        # reference values from the EXISTING entry point: error at the current values and after one damped trial step
        want0 = e.error(0)
        e.linearize()
        e.factorize(float(tag.split("_")[-2]), tag.endswith("_1"))
        e.backsolve()
        e.retract()
        want1 = e.error(1)
        assert want0 != want1
        bk = _Buckets(problem, torch, dev)
        up = lambda t, which: None if problem.n_vars(t) == 0 else torch.from_numpy(e.variables(t, which)).to(dev)
        cur = [up(t, 0) for t in range(4)]
        trial = [up(t, 1) for t in range(4)]
        assert lib.gtb_noalloc_setup(0) == capi.OK, lib.gtb_noalloc_last_error().decode()
        stream = torch.cuda.current_stream().cuda_stream
        torch.cuda.synchronize()
        mem_before = torch.cuda.memory_allocated()
        got0 = bk.total(lib, cur, stream)       # first call
        got1 = bk.total(lib, trial, stream)     # second call, other inputs: the shared scratch holds the first call's partials
        got0_again = bk.total(lib, cur, stream)
        assert torch.cuda.memory_allocated() == mem_before
        assert abs(got0 - want0) <= REL_TOL * abs(want0), (got0, want0)
        assert abs(got1 - want1) <= REL_TOL * abs(want0), (got1, want1)
        assert got0_again == got0   # fixed-order sums: bit-identical on repeat
    finally:
        e.close()


@pytest.mark.gpu
def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    lib = capi.lib()
    out = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    dummy = C.c_void_p(out.data_ptr())
    header = open(capi.HEADER).read()
    limit = 1 << 22
    assert "#define GTB_NOALLOC_MAX_FACTORS (1 << 22)" in header
    st = lib.gtb_bucket_error_noalloc(0, 2, limit + 1, dummy, dummy, dummy, dummy, dummy, None, None, dummy, None)
    assert st == capi.ERR_INVALID
    assert b"GTB_NOALLOC_MAX_FACTORS" in lib.gtb_noalloc_last_error()
    # an empty bucket is fine and gives zero
    out.fill_(7.0)
    st = lib.gtb_bucket_error_noalloc(0, 2, 0, None, None, None, None, None, None, None, dummy, None)
    assert st == capi.OK
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == 0.0


def test_noalloc_argument_checks_need_no_device():
    lib = capi.lib()
    assert lib.gtb_bucket_error_noalloc(0, 99, 1, None, None, None, None, None, None, None, None, None) == capi.ERR_INVALID
    assert lib.gtb_bucket_error_noalloc(0, 2, -1, None, None, None, None, None, None, None, None, None) == capi.ERR_INVALID
    assert lib.gtb_bucket_error_noalloc(0, 2, (1 << 22) + 1, None, None, None, None, None, None, None, None, None) == capi.ERR_INVALID
    assert lib.gtb_bucket_error_noalloc(0, 2, 4, None, None, None, None, None, None, None, None, None) == capi.ERR_INVALID
