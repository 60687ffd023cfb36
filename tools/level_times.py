"""Per-level device time of the large-front chain of one factorisation (per-launch CUDA events): which levels of the
elimination tree cost what.  usage: level_times.py [workload] [ordering]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from gtsam_petercdev_b200.engine import Engine
wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
ordn = sys.argv[2] if len(sys.argv) > 2 else "metis"
p = bench.make_problem(wl)
e = Engine(p, device=0)
e.set_ordering(bench.ordering_for(p, wl, ordn))
e.set_stage_timing(False)
for _ in range(3):
    e.linearize(); e.factorize(1e-5); e.backsolve()
e.linearize()
e.profile(True); e.factorize(1e-5)
names, ms = e.profile_read_raw()
e.profile(False)
# a level of large fronts = [large_children?] front_step (schur_step | front_step)*; a new level starts at a front_step that
# follows a schur_step (or the first one)
lv, cur, prev = [], None, None
pre = 0.0
for n, t in zip(names, ms):
    if n in ("front_step", "schur_step", "large_children"):
        if cur is None or (n in ("front_step", "large_children") and prev == "schur_step"):
            cur = {"front_step": 0.0, "schur_step": 0.0, "large_children": 0.0, "nf": 0, "ns": 0}
            lv.append(cur)
        cur[n] += t
        cur["nf" if n == "front_step" else "ns"] += n != "large_children"
        prev = n
    elif cur is None:
        pre += t
print("before the chain: %.3f ms" % pre)
tot = 0
for i, l in enumerate(lv):
    s = l["front_step"] + l["schur_step"] + l["large_children"]
    tot += s
    print("chain level %2d: %2d panel launches %.3f ms, %2d schur launches %.3f ms, gather %.3f ms  -> %.3f ms (cum %.3f)" % (
        i + 1, l["nf"], l["front_step"], l["ns"], l["schur_step"], l["large_children"], s, tot))
