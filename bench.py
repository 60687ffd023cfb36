#!/usr/bin/env python
"""bench.py -- LM iterations/s of the batch nonlinear least-squares hot path (BASELINE.json metric).

A "step" is one Levenberg-Marquardt inner iteration on one synthetic problem, exactly the stages the reference runs
per lambda trial (nonlinear/LevenbergMarquardtOptimizer.cpp:121-304): linearize -> damped multifrontal Cholesky ->
back-substitution -> linear.error(0), linear.error(delta) -> retract -> nonlinear error.  The trial is never accepted,
so every step does identical work on identical inputs (the same loop as `ref_tool mode=time`, the CPU baseline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c1] [--ordering metis|colamd|schur|natural]
    python bench.py --impl reference ...     # the unmodified reference (oracle/_ref) on the host cores

value     : device-resident throughput (inputs already in HBM), CUDA events on the engine stream, max over ranks
e2e       : the same step driven through the C ABI with HOST buffers: variables re-uploaded every step, the new error
            read back
roofline  : the kernel class with the largest share of the step, measured live with per-launch CUDA events
cpu_baseline : oracle/_ref/ref_tool (unmodified GTSAM 4.3a0, serial: TBB is not installed) on rank 0
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c3": "C3 Ladybug-scale synthetic BA (1,723 cams, 156,000 points, 679,032 GeneralSFMFactor<Cal3Bundler> + 2 priors)",
    "c2": "C2 sphere2500-shape Pose3 graph (2,500 Pose3, 4,949 BetweenFactor<Pose3> + prior)",
    "c1": "C1 M3500-shape Pose2 graph (3,500 Pose2, 5,450 BetweenFactor<Pose2> + prior)",
    "c4": "C4 Venice-scale synthetic BA (1,778 cams, ~993,000 points, ~5M GeneralSFMFactor<Cal3Bundler> + 2 priors)",
    "c2_real": "C2 on the shipped examples/Data/sphere2500.txt (2,500 Pose3, 4,949 BetweenFactor<Pose3> + prior; dead-reckoning initial values)",
    "c5": "C5 ISAM2 on City10000 (examples/ISAM2_City10000.cpp schedule: one update per measurement line, relinearizeThreshold 0.01, relinearizeSkip 1)",
}
CITY_FULL = os.path.join(ROOT, "tests", "golden", "data", "T1_city10000_04.txt")
DGEMM_FP64_TFLOPS = 35.46   # cuBLAS fp64 8192^3 measured on this pool's B200 (tools/probe_box.py, profiles/probe_r01.json)


def make_problem(workload):
    from gtsam_petercdev_b200 import datasets
    if workload == "c3":
        return datasets.ladybug_shape()
    if workload == "c2":
        return datasets.sphere_shape()
    if workload == "c1":
        return datasets.m3500_shape()
    if workload == "c4":
        return datasets.venice_shape()
    if workload == "c2_real":
        return datasets.sphere2500_file()
    raise SystemExit(f"unknown workload {workload}")


def ordering_for(problem, workload, name):
    """Ordering keys, or a builtin name.  COLAMD / METIS orderings are the reference's own (computed once by
    oracle/_ref/ref_tool, committed as permutations under tests/golden/): ordering is an input of the path."""
    if name in ("schur", "natural"):
        return name
    if name == "nested_dissection":
        # the library's own ordering (gtb_set_ordering_builtin(2)) as keys, so that the reference arm eliminates in the same order:
        # computed on a host-only engine, read back as the frontal keys of the cliques in post-order (an equivalent order)
        from gtsam_petercdev_b200.engine import Engine
        e = Engine(problem, device=-1)
        e.set_ordering("nested_dissection")
        pl = e.plan()
        starts = np.cumsum(pl["clique_nkeys"]) - pl["clique_nkeys"]
        keys = np.concatenate([pl["clique_keys"][s0:s0 + f] for s0, f in zip(starts, pl["clique_nfrontal"])])
        e.close()
        return keys.astype(np.uint64)
    if workload == "c2_real":
        return problem.variables()[0][np.load(os.path.join(ROOT, "tests", "golden", "fullsize__c2_real_metis.npz"))["ordering_" + name]]
    f = os.path.join(ROOT, "tests", "golden", {"c3": "ladybug_shape", "c2": "sphere_shape", "c1": "m3500_shape", "c4": "venice_shape"}[workload] + "_orderings.npz")
    perm = np.load(f)[name]
    return problem.variables()[0][perm]


def clocks_start(path):
    q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        return subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "20"],
                                stdout=open(path, "w"), stderr=subprocess.DEVNULL)
    except Exception:
        return None


def clocks_stop(proc, path, device):
    out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    if proc is None:
        return out
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()
    sm, mx, reasons = [], [], set()
    try:
        for line in open(path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9 or not f[0].isdigit() or int(f[0]) != device:
                continue
            sm.append(float(f[1]))
            mx.append(float(f[2]))
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
    except Exception:
        pass
    if sm:
        # "under load": the upper half of the samples (idle samples before/after the timed region pull the median down)
        s = sorted(sm)
        out["sm_mhz"] = float(np.median(s[len(s) // 2:]))
        out["sm_max_mhz"] = float(max(mx))
    out["reasons"] = sorted(reasons)
    return out


_JSON_OUT = None


def emit(obj):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(obj) + "\n")
    out.flush()


def bench_config(args):
    """The `config` object, identical in both arms (the driver compares them key by key)."""
    return {"workload": WORKLOADS[args.workload], "ordering": args.ordering, "lambda": args.lam,
            "l2": "inputs larger than L2 (no flush needed)" if args.workload in ("c3", "c4") else "inputs fit L2 (pose graph of the named size)"}


def source_digest():
    """SHA-256 over the device code (the .cuh files: every kernel lives there): an ncu traffic file is only quoted for the kernels
    it was taken from.  Host-side files (engine.cu: tables, scheduler, C ABI; the .cpp files) do not change what a launch reads."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "gtsam_petercdev_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith(".cuh"):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def dgemm_probe(torch, n=8192, reps=5):
    """FP64 matrix-multiply rate of THIS box, measured live (cuBLAS DGEMM n^3 through torch, best of `reps`): the
    denominator of the tensor-bound roofline lines (MEASURED_PEAKS.json carries no FP64 figure)."""
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(a, b)
    best = 1e30
    for _ in range(reps):
        s, f = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        torch.matmul(a, b)
        f.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(f))
    del a, b
    return 2.0 * n ** 3 / (best / 1e3) / 1e12


def run_reference(args, rank, world):
    """The reference's own CPU implementation of the step (oracle/_ref/ref_tool mode=time: unmodified GTSAM 4.3a0
    linked from oracle/_ref/libgtsam_ref.so), rank 0 only."""
    if rank != 0:
        return
    tool = os.path.join(ROOT, "oracle", "_ref", "ref_tool")
    base = {"impl": "reference", "metric": "LM iter/s (linearize+multifrontal Cholesky+backsub)", "unit": "iter/s",
            "n_gpus": args.gpus, "higher_is_better": True, "scaling": "strong" if args.parallelism == "tree" else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": bench_config(args)}
    if not os.path.exists(tool):
        emit({"impl": "reference", "unavailable": "oracle/_ref/ref_tool not built (needs /root/reference)"})
        return
    p = make_problem(args.workload)
    o = ordering_for(p, args.workload, args.ordering)
    tmp = tempfile.mkdtemp()
    if not isinstance(o, str):
        p.a["ordering"] = np.ascontiguousarray(o, dtype=np.uint64)
    pf, of = os.path.join(tmp, "p.gtb"), os.path.join(tmp, "o.gtb")
    p.save(pf)
    # bounded sample: at ~10 s per C3 step the run is capped to what fits in a few minutes
    est = {"c3": 11.0, "c2": 0.2, "c2_real": 0.2, "c1": 0.1, "c4": 120.0}[args.workload]
    steps = max(1, min(args.steps, int(240 / est)))
    warm = 1 if (args.warmup + steps) * est > 240 else args.warmup
    sample = f"{steps} full LM inner iterations of the whole workload after {warm} warm-up (asked {args.steps}/{args.warmup})"
    from gtsam_petercdev_b200 import gtbio
    oname = o if isinstance(o, str) else "file"
    t0 = time.time()
    subprocess.run([tool, pf, of, "mode=time", f"ordering={oname}", f"iters={steps}", f"warm={warm}", f"lambda={args.lam!r}"],
                   check=True)
    wall = time.time() - t0
    r = gtbio.load(of)
    t_iter = float(np.mean(r["t_iter"]))
    base.update({"value": 1.0 / t_iter, "steps": steps, "warmup": warm, "ms_per_step": 1e3 * t_iter,
                 "cpu_baseline": {"value": 1.0 / t_iter, "unit": "iter/s", "cores": 1, "kind": "reference", "sample": sample,
                                  "host_cores": os.cpu_count(), "threads_note": "TBB not installed: the reference runs serial",
                                  "stage_ms": {k: 1e3 * float(np.mean(r[k])) for k in ("t_linearize", "t_eliminate", "t_backsub")},
                                  "wall_s": wall},
                 "e2e": {"value": 1.0 / t_iter, "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    emit(base)


def isam2_reference(lines, loops_hint=None):
    """The reference's own gtsam::ISAM2 on the first `lines` lines (oracle/_ref/ref_tool mode=isam2), seconds per update."""
    from gtsam_petercdev_b200 import gtbio
    tool = os.path.join(ROOT, "oracle", "_ref", "ref_tool")
    if not os.path.exists(tool):
        return None
    tmp = tempfile.mkdtemp()
    head = os.path.join(tmp, "head.txt")
    with open(CITY_FULL) as f, open(head, "w") as g:
        for i, line in enumerate(f):
            if i >= lines:
                break
            g.write(line)
    out = os.path.join(tmp, "o.gtb")
    subprocess.run([tool, head, out, "mode=isam2", "every=1000000"], check=True)
    return np.asarray(gtbio.load(out)["update_seconds"], dtype=np.float64)


def latency_stats(ms):
    ms = np.asarray(ms, dtype=np.float64)
    return {"mean": float(ms.mean()), "median": float(np.median(ms)), "p99": float(np.percentile(ms, 99)), "max": float(ms.max()),
            "last_100_mean": float(ms[-100:].mean())}


def run_isam2(args, reference_arm):
    """--workload c5: per-update latency of the ISAM2 layer (gtb_isam2_update + read-back of the estimate, host buffers, through
    the C ABI) over the first --updates lines of City10000, beside the reference's gtsam::ISAM2 on the same lines on this
    host.  A step = one update; value = updates per second over the whole run (wall clock: the update is host-driven)."""
    lines = args.updates
    cfg = {"workload": WORKLOADS["c5"], "updates": lines, "ordering": "built-in nested dissection, recomputed every 8 updates (device arm); "
           "constrained COLAMD per update (reference arm)"}
    base = {"metric": "ISAM2 updates/s (City10000 schedule)", "unit": "updates/s", "n_gpus": 1, "higher_is_better": True, "scaling": "n/a",
            "vs_baseline": None, "dtype": "f64", "data": "the reference's examples/Data/T1_city10000_04.txt (first %d lines)" % lines,
            "config": cfg, "steps": lines, "warmup": 0}
    if reference_arm:
        sec = isam2_reference(lines)
        if sec is None:
            emit({"impl": "reference", "unavailable": "oracle/_ref/ref_tool not built (needs /root/reference)"})
            return
        ms = 1e3 * sec
        base.update({"impl": "reference", "value": float(len(ms) / sec.sum()), "ms_per_step": float(ms.mean()), "latency_ms": latency_stats(ms),
                     "cpu_baseline": {"value": float(len(ms) / sec.sum()), "unit": "updates/s", "cores": 1, "kind": "reference",
                                      "sample": "all %d updates, gtsam::ISAM2 (ref_tool mode=isam2)" % len(ms)},
                     "e2e": {"value": float(len(ms) / sec.sum()), "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                     "gpu_launches": 0})
        emit(base)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    from gtsam_petercdev_b200.isam2 import ISAM2, drive_city10000, read_city10000
    rows = read_city10000(CITY_FULL)[:lines]
    warm = ISAM2(0.01, 1)                      # CUDA context, module load, allocator warm-up: untimed
    drive_city10000(warm, rows[:50])
    warm.close()
    sampler = clocks_start("/tmp/gtb_clocks_isam2.csv")
    isam = ISAM2(0.01, 1)
    stamps = [time.perf_counter()]
    est, row_of, results = drive_city10000(isam, rows, lambda u, e, r: stamps.append(time.perf_counter()))
    clocks = clocks_stop(sampler, "/tmp/gtb_clocks_isam2.csv", 0)
    wall_ms = 1e3 * np.diff(np.array(stamps))
    total_s = stamps[-1] - stamps[0]
    lib = isam.lib
    launches = int(lib.gtb_kernel_launches(lib.gtb_isam2_engine(isam.h)))
    n = len(rows)
    h2d = int(np.mean([r["n_variables"] * 3 * 8 + r["n_factors"] * (2 * 8 + 3 * 8 + 4 + 9 * 8) for r in results]))
    d2h = int(np.mean([r["n_variables"] * 4 * 8 for r in results]))
    base.update({"value": n / total_s, "ms_per_step": float(wall_ms.mean()), "latency_ms": latency_stats(wall_ms),
                 "host_plan_ms": latency_stats([r["plan_ms"] for r in results]), "device_solve_ms": latency_stats([r["solve_ms"] for r in results]),
                 "e2e": {"value": n / total_s, "unit": "updates/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                 "gpu_launches": launches, "gpu_launches_note": "kernels of the LAST update (every update builds a fresh engine)",
                 "clocks": clocks, "final": {"poses": results[-1]["n_variables"], "factors": results[-1]["n_factors"], "cliques": results[-1]["n_cliques"],
                                             "relinearized_variables_total": int(sum(r["variables_relinearized"] for r in results))},
                 "roofline": {"bound": "launch latency + host planning", "note": "an update of a <= 10k-pose graph is a few hundred kernels of microseconds "
                              "each plus a symbolic analysis on the host: neither HBM nor the FP64 pipe is the bound (see host_plan_ms / device_solve_ms)",
                              "achieved": None, "peak": None, "frac": None, "traffic": None}})
    if not args.no_cpu_baseline:
        sec = isam2_reference(lines)
        if sec is not None:
            base["cpu_baseline"] = {"value": float(len(sec) / sec.sum()), "unit": "updates/s", "cores": 1, "kind": "reference",
                                    "sample": "the same %d updates, gtsam::ISAM2 (ref_tool mode=isam2) on this host" % len(sec),
                                    "latency_ms": latency_stats(1e3 * sec)}
    emit(base)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--ordering", default="metis")
    ap.add_argument("--lam", type=float, default=1e-5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--updates", type=int, default=1264, help="--workload c5: number of City10000 lines (1264 = the first 200 loop closures)")
    ap.add_argument("--parallelism", default="tree", choices=["tree", "replicas"],
                    help="N > 1: 'tree' = ONE problem, elimination tree cut across the GPUs (fan-in over NCCL, strong "
                         "scaling); 'replicas' = one independent problem per GPU (weak scaling, no collective)")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: everything else that libraries print there (NCCL's version banner, the
    # reference tool's progress) goes to stderr
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "c5":
        if rank == 0:
            run_isam2(args, args.impl == "reference")
        return
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from gtsam_petercdev_b200.engine import Engine

    # ---- set-up (untimed): problem, upload, symbolic plan ----------------------------------------------------
    p = make_problem(args.workload)
    t0 = time.time()
    e = Engine(p, device=local)
    t_upload = time.time() - t0
    t0 = time.time()
    tree = world > 1 and args.parallelism == "tree"
    if tree:
        from gtsam_petercdev_b200.distributed import torch_comm
        e.set_distributed(rank, world, torch_comm(local))
    e.set_ordering(ordering_for(p, args.workload, args.ordering))
    t_plan = time.time() - t0
    plan = e.plan()
    st = plan["stats"]
    stream = torch.cuda.ExternalStream(e.stream(), device=torch.device("cuda", local))
    e.set_stage_timing(False)
    lam = args.lam

    def step():
        # one LM trial = linearize + what tryLambda does (LevenbergMarquardtOptimizer.cpp:121-270): damped multifrontal
        # Cholesky, back-substitution, both linear errors, retract, nonlinear error; gtb_trial enqueues them back to back
        # and synchronises once (the three scalars and the failure flag come back together)
        e.linearize()
        return e.trial(lam)[2]

    cam, pt = p.a.get("cam"), p.a.get("point")
    # the step's inputs live in page-locked host memory (the contract's "copy from pinned host memory"): the library then copies
    # straight from these buffers, without a staging pass
    pinned = [(t, torch.from_numpy(np.ascontiguousarray(p.a[n])).pin_memory()) for t, n in enumerate(["pose2", "pose3", "cam", "point"]) if n in p.a]
    host_state = [(t, x.numpy()) for t, x in pinned]
    h2d = int(sum(a.nbytes for _, a in host_state))

    def step_e2e():
        for t, a in host_state:
            e.set_state(t, a)       # host -> device copy of the step's inputs through the C ABI
        return step()               # the new error comes back to the host (8 bytes) inside error()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        err1 = step()
    clk_path = os.path.join(tempfile.gettempdir(), f"gtb_clocks_{rank}.csv")
    clk = clocks_start(clk_path) if rank == 0 else None
    # nvidia-smi needs a few hundred ms before its first sample: keep the GPU under the same load (untimed steps) until the
    # sampler is live, so that even a short timed region is covered by samples taken during it.  With several ranks every
    # rank runs the same fixed number of extra steps (a step contains collectives).
    if world > 1:
        for _ in range(150):
            step()
    elif clk is not None:
        t_wait = time.time()
        while time.time() - t_wait < 3.0:
            step()
            try:
                if os.path.getsize(clk_path) > 0:
                    break
            except OSError:
                pass

    def timed(fn, n):
        barrier()
        l0 = e.kernel_launches()
        s, f = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(stream)
        for _ in range(n):
            fn()
        f.record(stream)
        barrier()
        ms = s.elapsed_time(f)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, e.kernel_launches() - l0

    ms_dev, launches = timed(step, args.steps)
    for _ in range(2):
        step_e2e()          # untimed: first use allocates the engine's pinned staging buffers
    ms_e2e, _ = timed(step_e2e, args.steps)
    clocks = clocks_stop(clk, clk_path, local) if rank == 0 else None

    # ---- per-kernel-class breakdown, measured live with per-launch CUDA events ---------------------------------
    e.profile(True)
    nprof = min(args.steps, 3)
    for _ in range(nprof):
        step()
    prof = e.profile_read()
    e.profile(False)
    e.set_stage_timing(True)
    step()
    stage = e.stage_ms()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    per_step = {k: (v[0] / nprof, v[1] // nprof) for k, v in prof.items()}
    total_k = sum(v[0] for v in per_step.values())
    share = {k: v[0] / total_k for k, v in per_step.items()}
    hbm = 6538.6
    dgemm = dgemm_probe(torch)
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        hbm = float(peaks["hbm_gbs"])
        peak_src = (f"HBM: MEASURED_PEAKS.json hbm_gbs ({peaks.get('when', '?')}, {peaks.get('gpu_name', '?')}); FP64: cuBLAS DGEMM 8192^3 "
                    f"measured live in this run on this box = {dgemm:.2f} TFLOP/s (MEASURED_PEAKS.json has no FP64 figure; "
                    f"tcgen05 has no f64 kind, the FP64 tensor instruction is mma.sync m8n8k4)")
    except Exception:
        hbm, peak_src = 6650.0, f"HBM: fallback of the profiling recipe; FP64: cuBLAS DGEMM 8192^3 measured live = {dgemm:.2f} TFLOP/s"
    # algorithmic work of THIS rank (rank 0): with several ranks only the cliques rank 0 factors count (its subtrees and
    # the shared top of the tree), computed from the reference clique layout and the ownership map
    flops_panel, flops_schur = st["flops_large"] - st["flops_syrk_schur"], st["flops_syrk_schur"]
    if tree:
        own = e.partition()["owner"]
        nf, nk = plan["clique_nfrontal"], plan["clique_nkeys"]
        dims = dict(zip(e.var_keys.tolist(), e.var_dim.tolist()))
        off, fp, fs = 0, 0.0, 0.0
        keys = plan["clique_keys"].tolist()
        for c in range(len(nf)):
            kk = keys[off:off + int(nk[c])]
            off += int(nk[c])
            if own[c] > 0:
                continue
            F = float(sum(dims[k] for k in kk[:int(nf[c])]))
            S = float(sum(dims[k] for k in kk[int(nf[c]):]))
            if F + S + 1 <= 128 or (F <= 12 and F + S + 1 <= 208):
                continue          # small fronts are not panel / Schur work
            s1 = S + 1
            fs += F * s1 * (s1 + 1)
            fp += F ** 3 / 3 + F * F * s1
        flops_panel, flops_schur = fp, fs
    # algorithmic work per step of each candidate class
    work = {
        "front_step": ("tensor", flops_panel / 1e12, dgemm, "TFLOP/s"),
        "schur_step": ("tensor", flops_schur / 1e12, dgemm, "TFLOP/s"),
        "small_front": ("hbm", st["bytes_dense_small"] / 1e9, hbm, "GB/s"),
        "thin_front": ("hbm", st["bytes_thin"] / 1e9, hbm, "GB/s"),
        "point_front": ("hbm", st["bytes_point"] / 1e9, hbm, "GB/s"),
        "linearize": ("hbm", st["bytes_linearize"] / 1e9, hbm, "GB/s"),
        "backsolve": ("hbm", st["bytes_backsolve"] / 1e9, hbm, "GB/s"),
    }
    dom = max((k for k in per_step if k in work), key=lambda k: per_step[k][0])
    bound, amount, peak, unit = work[dom]
    achieved = amount / (per_step[dom][0] / 1e3)
    # DRAM bytes per launch of the dominant class, from the committed ncu pass over every launch of this same command
    # (quoted only when the file was taken from THIS build of the kernels and on one GPU: it records the digest of the sources)
    traffic, traffic_src = None, None
    tf = os.path.join(ROOT, "profiles", "r02", "ncu_traffic_c3_metis.json")
    if args.workload == "c3" and args.ordering == "metis" and world == 1 and os.path.exists(tf):
        tj = json.load(open(tf))
        if tj.get("_source_digest") != source_digest():
            traffic_src = "profiles/r02/ncu_traffic_c3_metis.json was taken from another build of the kernels: not quoted"
        elif dom in tj:
            traffic, traffic_src = tj[dom]["traffic_bytes_per_launch"], "profiles/r02/ncu_traffic_c3_metis.json (" + tj["_source"] + ")"
    roofline = {"kernel": dom, "bound": bound, "achieved": achieved, "peak": peak, "unit": unit, "frac": achieved / peak,
                "traffic": traffic, "traffic_unit": "bytes per launch (dram read + write)", "traffic_source": traffic_src, "peak_source": peak_src, "share_of_step": share[dom],
                "launches_per_step": per_step[dom][1], "ms_per_step": per_step[dom][0],
                "all_classes_ms_per_step": {k: round(v[0], 4) for k, v in sorted(per_step.items(), key=lambda kv: -kv[1][0])},
                "other": {k: {"achieved": work[k][1] / (per_step[k][0] / 1e3), "unit": work[k][3], "frac": work[k][1] / (per_step[k][0] / 1e3) / work[k][2]}
                          for k in work if k in per_step and k != dom}}

    units = 1 if (tree or world == 1) else world     # problems advanced per step by the whole job
    if world == 1:
        par = "single GPU"
    elif tree:
        pt = e.partition()
        par = (f"one problem over {world} GPUs: elimination tree cut into per-rank subtrees, {pt['n_top']} top fronts summed onto "
               f"rank 0 over NCCL ({pt['reduced_doubles'] * 8 / 1e6:.0f} MB per step), delta broadcast + all-reduce; "
               f"modelled cost shares top={pt['top_share']:.2f} ranks={[round(float(x), 2) for x in pt['rank_share'][:world]]}")
    else:
        par = f"{world} independent replicas (one problem per GPU, no collective)"
    line = {
        "metric": "LM iter/s (linearize+multifrontal Cholesky+backsub)", "value": units * args.steps / (ms_dev / 1e3),
        "unit": "iter/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        # ONE problem whatever N: the tree mode is strong scaling (also on the N = 1 line of a scaling run)
        "scaling": "strong" if args.parallelism == "tree" else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": bench_config(args),
        "details": {"parallelism": par, "ordering_source": "the reference's own Ordering::Metis / Colamd, precomputed (tests/golden/*_orderings.npz)",
                    "working_set": "fronts %.1f GB, Jacobians %.0f MB" % (st["arena_doubles"] * 8 / 1e9, st["bytes_linearize"] / 1e6),
                    "cliques": int(len(plan["clique_parent"])), "levels_of_the_numeric_plan": int(st["levels"]), "max_front": int(st["max_front"]),
                    "elimination_gflop": st["flops"] / 1e9, "setup_s": {"upload": t_upload, "symbolic_plan": t_plan}},
        # the step's result that comes back to the host: the failure flag, both linear errors and the new graph error
        # (9 doubles); delta and the retracted values stay on the device, as in an LM trial of the reference's loop
        "e2e": {"value": units * args.steps / (ms_e2e / 1e3), "unit": "iter/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8 * 9},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "stage_ms": stage,
        "fp64_tflops_elimination": st["flops"] / 1e12 / (stage["factorize"] / 1e3),
        "trial_error": err1,
    }
    # ---- CPU baseline: the unmodified reference on the host, bounded sample ------------------------------------
    tool = os.path.join(ROOT, "oracle", "_ref", "ref_tool")
    if not args.no_cpu_baseline and world == 1 and os.path.exists(tool):
        from gtsam_petercdev_b200 import gtbio
        tmp = tempfile.mkdtemp()
        o = ordering_for(p, args.workload, args.ordering)
        if not isinstance(o, str):
            p.a["ordering"] = np.ascontiguousarray(o, dtype=np.uint64)
        pf, of = os.path.join(tmp, "p.gtb"), os.path.join(tmp, "o.gtb")
        p.save(pf)
        iters = 1 if args.workload in ("c3", "c4") else 5
        subprocess.run([tool, pf, of, "mode=time", f"ordering={o if isinstance(o, str) else 'file'}", f"iters={iters}", "warm=1",
                        f"lambda={lam!r}"], check=True)
        r = gtbio.load(of)
        t_iter = float(np.mean(r["t_iter"]))
        line["cpu_baseline"] = {"value": 1.0 / t_iter, "unit": "iter/s", "cores": 1, "kind": "reference",
                                "sample": f"{iters} full LM inner iteration(s) of the same workload after 1 warm-up, oracle/_ref/ref_tool mode=time",
                                "host_cores": os.cpu_count(), "threads_note": "TBB not installed: the reference runs serial",
                                "stage_ms": {k: 1e3 * float(np.mean(r[k])) for k in ("t_linearize", "t_eliminate", "t_backsub")},
                                "trial_error": float(r["trial_error"][0]) if "trial_error" in r else None}
    else:
        line["cpu_baseline"] = {"value": None, "unit": "iter/s", "cores": 0, "kind": "reference",
                                "sample": "skipped (multi-GPU run, --no-cpu-baseline, or oracle/_ref missing)"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
