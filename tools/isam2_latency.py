"""Per-update latency of gtb_isam2_update on the City10000 schedule (tests/test_isam2_gpu.py drives the same loop): wall time per
update with its split into host planning and device solve, by hundreds of updates."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import test_isam2_gpu as T
from gtsam_petercdev_b200.isam2 import ISAM2

rows = T.read_city(sys.argv[1] if len(sys.argv) > 1 else T.HEAD)
if len(sys.argv) > 2:
    rows = rows[:int(sys.argv[2])]
isam = ISAM2(0.01, 1)
stamps = [time.perf_counter()]
est, row_of, results = T.drive(isam, rows, lambda u, e, r: stamps.append(time.perf_counter()))
wall = np.diff(np.array(stamps)) * 1e3
plan = np.array([r["plan_ms"] for r in results])
solve = np.array([r["solve_ms"] for r in results])
for a in range(0, len(rows), max(100, len(rows) // 12)):
    b = min(len(rows), a + max(100, len(rows) // 12))
    print("updates %5d-%5d  poses %5d  wall %7.3f ms  plan %7.3f  solve %7.3f  cliques %d" % (
        a, b, results[b - 1]["n_variables"], wall[a:b].mean(), plan[a:b].mean(), solve[a:b].mean(), results[b - 1]["n_cliques"]))
print("all: wall mean %.3f median %.3f p99 %.3f ms; plan mean %.3f; solve mean %.3f" % (
    wall.mean(), np.median(wall), np.percentile(wall, 99), plan.mean(), solve.mean()))
