#!/bin/bash
# Trimmed round-end pass: GPU tests, smoke, default bench (with the CPU baseline), pose-graph lines, launch trace.
# FULL=1 adds the COLAMD lines and the ncu launch list with DRAM bytes (about 4 more minutes).
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
python bench.py > gpurun_out/bench_c3_full.json 2> gpurun_out/bench_c3_full.err
for cfg in "c1 metis" "c2 metis" ${FULL:+"c1 colamd" "c2 colamd"}; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps 10 > gpurun_out/bench_$1_$2.json 2> gpurun_out/bench_$1_$2.err
done
GTB_DEBUG=8 python tools/launch_trace.py c3 metis > gpurun_out/launch_trace_c3_metis.txt 2>&1
if [ -n "$FULL" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv \
    --log-file gpurun_out/ncu_launches_c3_metis.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
  gzip -f gpurun_out/ncu_launches_c3_metis.csv
fi
cat gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log
for f in gpurun_out/bench_c*_*.json; do python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"), d.get("cpu_baseline",{}).get("value"))
except Exception as e: print(sys.argv[1], "ERR", e)
P
done
