"""Attribute ncu warp-stall samples (source page, SASS view, csv) to source lines through nvdisasm -g line markers.
usage: ncu_lines.py sass.csv dis.txt kernel_substring [top]"""
import csv
import re
import sys
from collections import defaultdict

sass_csv, dis, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
# nvdisasm: instructions of the kernel in order with the current line marker
lines = open(dis).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and kern in l)
cur, seq = None, []
for l in lines[start + 1:]:
    if l.startswith(".text.") or l.startswith("//--------------------- .text"):
        if seq:
            break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        seq.append(cur)
rows = [r for r in csv.reader(open(sass_csv))]
ends = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
rows = rows[ends[0]:ends[1]] if len(ends) > 1 else rows      # first captured launch only
hdr = rows[1]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr)]
print(f"# {len(body)} SASS instructions in ncu, {len(seq)} in nvdisasm")
agg = defaultdict(lambda: [0, 0, defaultdict(int)])
for k, r in enumerate(body):
    key = seq[k] if k < len(seq) and len(seq) == len(body) else ("?", 0)
    a = agg[key]
    a[0] += int(r[isamp] or 0)
    a[1] += int(r[iex] or 0)
    for i, h in stall_cols:
        a[2][h] += int(r[i] or 0)
tot = sum(a[0] for a in agg.values())
src = {}
for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    if f not in src:
        try:
            src[f] = open("/root/repo/gtsam_petercdev_b200/csrc/" + f).read().split("\n")
        except OSError:
            src[f] = []
    text = src[f][ln - 1].strip()[:90] if 0 < ln <= len(src[f]) else ""
    st = ", ".join(f"{h[6:]}={v}" for h, v in sorted(a[2].items(), key=lambda kv: -kv[1])[:3] if v)
    print(f"{a[0] / tot:6.3f} {a[0]:7d} smp {a[1]:9d} inst  {f}:{ln:<5d} {text}   [{st}]")
