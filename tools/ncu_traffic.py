"""DRAM traffic per launch by kernel class from an `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,... --csv` log.
Writes the json that bench.py reads for roofline.traffic.   usage: ncu_traffic.py traffic.csv out.json"""
import csv
import json
import re
import sys
from collections import defaultdict

CLASS = {"front_step_kernel": "front_step", "schur_step_kernel": "schur_step", "thin_front_kernel": "thin_front", "point_front_kernel": "point_front", "point_rows_kernel": "point_front",
         "point_pair_kernel": "point_schur", "point_diag_kernel": "point_schur",
         "small_front_kernel": "small_front", "large_children_kernel": "large_children", "large_zero_kernel": "large_zero",
         "large_factors_kernel": "large_factors", "backsolve_kernel": "backsolve", "backsolve_gemv_kernel": "backsolve_gemv",
         "linearize_kernel": "linearize", "error_kernel": "error", "linear_error_kernel": "linear_error",
         "backsolve_points_kernel": "backsolve", "pair_reduce_kernel": "point_schur", "extend_add_kernel": "large_children"}
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main(path, out):
    lines = [l for l in open(path) if l.startswith('"')]
    per = defaultdict(lambda: defaultdict(float))   # launch id -> metric -> value
    name = {}
    for r in csv.DictReader(lines):
        m = r["Metric Name"]
        if not m.startswith("dram__bytes"):
            continue
        v = float(r["Metric Value"].replace(",", "")) * UNIT.get(r["Metric Unit"], 1.0)
        per[r["ID"]][m] += v
        name[r["ID"]] = re.sub(r"[<(].*", "", r["Kernel Name"]).replace("void ", "").split("::")[-1]
    agg = defaultdict(lambda: [0, 0.0, 0.0])
    for i, ms in per.items():
        k = CLASS.get(name[i], name[i])
        agg[k][0] += 1
        agg[k][1] += ms.get("dram__bytes_read.sum", 0.0)
        agg[k][2] += ms.get("dram__bytes_write.sum", 0.0)
    res = {k: {"launches": n, "dram_read_bytes_per_launch": r / n, "dram_write_bytes_per_launch": w / n,
               "traffic_bytes_per_launch": (r + w) / n} for k, (n, r, w) in agg.items()}
    import os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    res["_source_digest"] = bench.source_digest()   # bench.py quotes the file only for this build of the kernels
    res["_source"] = "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum over every launch of `bench.py --steps 2 --warmup 1` (C3, METIS)"
    json.dump(res, open(out, "w"), indent=1, sort_keys=True)
    for k, v in sorted(res.items()):
        if not k.startswith("_"):
            print(f"{k:18s} {v['launches']:5d} launches  {v['traffic_bytes_per_launch'] / 1e6:10.3f} MB / launch")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
