"""Per-launch device times of one LM step (CUDA events around every launch), grouped for reading: which kernel
class at which position of the level schedule costs what."""
import sys, os, json
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
def step():
    e.linearize(); e.factorize(1e-5); e.backsolve(); e.linear_errors(); e.retract(); return e.error(1)
for _ in range(3): step()
e.profile(True); step()
names, ms = e.profile_read_raw()
e.profile(False)
# compress consecutive launches of the same class
out, i = [], 0
while i < len(names):
    j = i
    while j < len(names) and names[j] == names[i]: j += 1
    seg = ms[i:j]
    out.append((names[i], j - i, float(seg.sum()), float(seg.max())))
    i = j
for r in out: print("%-16s n=%4d total=%8.3f ms  max=%7.3f ms" % r)
fs = ms[[k for k, n in enumerate(names) if n == "front_step"]]
print("front_step launches", len(fs), "sum", fs.sum(), "sorted top", np.sort(fs)[::-1][:20])

import ctypes
if os.environ.get("GTB_DEBUG") and int(os.environ["GTB_DEBUG"]) & 8:
    buf = (ctypes.c_longlong * 16)()
    e.lib.gtb_debug_read.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    e.lib.gtb_debug_read(e.h, buf)
    b = list(buf)
    print("role A cycle stamps (root, p=2, slice 0), warp 0, cycles from kernel entry:")
    print("  loads + sync               %7d" % (b[1] - b[0]))
    print("  D -= PtP + sync            %7d   (tile-pair indices %d, DMMA loop %d, write-back %d, barrier %d)" % (
        b[2] - b[1], b[4] - b[1], b[6] - b[4], b[7] - b[6], b[2] - b[7]))
    print("  first 8 x 8 block factored %7d" % (b[3] - b[2]))
    print("  64 x 64 Cholesky + inverses%7d" % (b[5] - b[2]))
    print("column warp 1: X update done at %d, released at %d, solve %d, store %d (cycles from kernel start)" % (
        b[8] - b[0], b[9] - b[0], b[10] - b[9], b[11] - b[10]))
    print("schur step of the root's last child: tile 0 %d cycles (%d ns), tile 77 %d cycles (%d ns)" % (b[12], b[13], b[14], b[15]))

if os.environ.get("GTB_TRACE_STEPS"):
    print("front_step launches in order (ms):")
    print(" ".join("%.3f" % x for x in fs))
    ss = ms[[k for k, n in enumerate(names) if n == "schur_step"]]
    print("schur_step launches in order (ms):")
    print(" ".join("%.3f" % x for x in ss))
