"""Marginal covariances of EVERY variable: device sweep (gtb_marginals_compute + one gather) next to the unmodified
reference's Marginals(graph, values) + marginalCovariance(key) per key (oracle/_ref/ref_tool mode=marginals) on the
host.  usage: marginals_bench.py c1|c2 [ordering]   -> one JSON line"""
import json, os, subprocess, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from gtsam_petercdev_b200 import gtbio
from gtsam_petercdev_b200.engine import Engine

wl = sys.argv[1] if len(sys.argv) > 1 else "c1"
ordn = sys.argv[2] if len(sys.argv) > 2 else "metis"
p = bench.make_problem(wl)
e = Engine(p, device=0)
e.set_ordering(bench.ordering_for(p, wl, ordn))
keys, _, _, dims = p.variables()
e.linearize(); e.factorize(0.0)
e.marginal_covariances(keys[:1])            # builds the tables, warm-up
ts = []
for _ in range(5):
    e.linearize(); e.factorize(0.0)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    covs = e.marginal_covariances(keys)     # sweep + gather + copy to the host
    torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
line = {"workload": bench.WORKLOADS[wl], "ordering": ordn, "variables": int(len(keys)),
        "device_all_marginals_ms": 1e3 * float(np.median(ts))}
tool = os.path.join(bench.ROOT, "oracle", "_ref", "ref_tool")
if os.path.exists(tool):
    tmp = tempfile.mkdtemp(); pf, of = os.path.join(tmp, "p.gtb"), os.path.join(tmp, "o.gtb")
    p.save(pf)
    subprocess.run([tool, pf, of, "mode=marginals"], check=True)
    r = gtbio.load(of)
    line["reference_all_marginals_ms"] = 1e3 * float(r["t_marginals"][0])
    o, worst = 0, 0.0
    for d, C in zip(r["marginal_dims"], covs):
        ref = r["marginal_cov"][o:o + d * d].reshape(d, d); o += d * d
        worst = max(worst, float(np.linalg.norm(C - ref) / np.linalg.norm(ref)))
    line["max_relative_block_error"] = worst
    line["speedup"] = line["reference_all_marginals_ms"] / line["device_all_marginals_ms"]
print(json.dumps(line))
