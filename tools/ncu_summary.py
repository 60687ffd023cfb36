"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: launches, total, share."""
import csv
import re
import sys
from collections import defaultdict


def main(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        ns = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("us", "usecond"):
            ns *= 1e3
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        rows.append((name, ns))
    # shares over this library's kernels only (bench.py also runs a cuBLAS DGEMM probe for the FP64 peak: not part of the step)
    ours = [(n, ns) for n, ns in rows if n.startswith("gtb::")]
    other = sum(ns for n, ns in rows if not n.startswith("gtb::"))
    rows = ours
    tot = sum(ns for _, ns in rows)
    agg = defaultdict(lambda: [0, 0.0, 0.0])
    for n, ns in rows:
        a = agg[n]
        a[0] += 1
        a[1] += ns
        a[2] = max(a[2], ns)
    print(f"# {path}: {len(rows)} launches of gtb:: kernels, {tot / 1e6:.3f} ms total (cold-cache, serialised); "
          f"other kernels in the process (cuBLAS DGEMM peak probe, torch RNG): {other / 1e6:.3f} ms, not part of the step")
    print(f"{'kernel':44s} {'launches':>8s} {'total ms':>10s} {'share':>7s} {'avg us':>9s} {'max us':>9s}")
    for n, (c, t, m) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:44s} {c:8d} {t / 1e6:10.3f} {t / tot:7.3f} {t / c / 1e3:9.2f} {m / 1e3:9.2f}")


if __name__ == "__main__":
    main(sys.argv[1])
