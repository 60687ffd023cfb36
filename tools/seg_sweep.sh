#!/bin/bash
# Segment-cap sweep of the camera-pair lists on C3 (one bench line per setting, no CPU baseline).
mkdir -p gpurun_out
for cfg in "32 32" "32 16" "32 8" "16 16"; do
  set -- $cfg
  GTB_PAIR_SEG=$1 GTB_DIAG_SEG=$2 python bench.py --steps 10 --no-cpu-baseline > gpurun_out/seg_$1_$2.json 2> gpurun_out/seg_$1_$2.err
done
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/seg_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); c = d["roofline"]["all_classes_ms_per_step"]
        print(f, round(d["ms_per_step"], 3), "point_schur", c["point_schur"], "trial_error", d.get("trial_error"))
    except Exception as ex:
        print(f, "ERR", ex)
P
