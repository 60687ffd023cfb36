"""The ISAM2 layer over the WHOLE City10000 file (20,687 updates) against the reference's own gtsam::ISAM2 run of the same file
(tests/golden/isam2__city10000_200loops.npz: full_final_stride50 = every 50th pose of its final estimate, full_counts): max
absolute difference of x, y, theta.  ~5.5 minutes on a B200."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gtsam_petercdev_b200.isam2 import ISAM2, drive_city10000, read_city10000

g = dict(np.load(os.path.join(ROOT, "tests", "golden", "isam2__city10000_200loops.npz")))
rows = read_city10000(os.path.join(ROOT, "tests", "golden", "data", "T1_city10000_04.txt"))
isam = ISAM2(0.01, 1)
est, row_of, results = drive_city10000(isam, rows)
n = len(row_of)
xyt = np.stack([est[:, 2], est[:, 3], np.arctan2(est[:, 1], est[:, 0])], axis=1)[[row_of[i] for i in range(n)]]
want = g["full_final_stride50"]
got = xyt[::50]
d = np.abs(got - want)
d[:, 2] = np.abs(np.angle(np.exp(1j * (got[:, 2] - want[:, 2]))))
updates, loops, nposes, nfactors = g["full_counts"].tolist()
out = {"updates": len(rows) + 1, "reference_updates": updates + 1, "poses": n, "reference_poses": nposes, "factors": results[-1]["n_factors"],
       "reference_factors": nfactors, "max_abs_diff_of_every_50th_pose": float(d.max()), "median_abs_diff": float(np.median(d)),
       "relinearized_variables_total": int(sum(r["variables_relinearized"] for r in results))}
print(json.dumps(out))
