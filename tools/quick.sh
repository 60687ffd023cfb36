#!/bin/bash
# Quick check on the box: GPU tests, then short bench lines without the CPU baseline.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/pytest_gpu.log
for cfg in "c3 metis" "c2 metis" "c1 metis" ${QUICK_EXTRA}; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps 10 --no-cpu-baseline > gpurun_out/q_$1_$2.json 2> gpurun_out/q_$1_$2.err
done
cat gpurun_out/pytest_gpu.log
python - <<'P'
import json, glob
for f in sorted(glob.glob("gpurun_out/q_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, round(d["value"], 1), round(d["ms_per_step"], 3), d["roofline"]["all_classes_ms_per_step"])
    except Exception as ex:
        print(f, "ERR", ex)
P
