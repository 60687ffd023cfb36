#!/bin/bash
# Round bench matrix on one B200: default bench (C3/METIS, with CPU baseline), other workloads/orderings, reference arm.
# Outputs land in gpurun_out/ (copy what should be judged into profiles/rNN/).
set -x
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_c3_full.json 2> gpurun_out/bench_c3_full.err
for cfg in "c3 colamd" "c3 schur" "c1 colamd" "c1 metis" "c2 colamd" "c2 metis"; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps 10 > gpurun_out/bench_$1_$2.json 2> gpurun_out/bench_$1_$2.err
done
python bench.py --workload c4 --ordering metis --steps 5 --no-cpu-baseline > gpurun_out/bench_c4_metis.json 2> gpurun_out/bench_c4_metis.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c3_ref.json 2> gpurun_out/bench_c3_ref.err
GTB_DEBUG=8 python tools/launch_trace.py c3 metis > gpurun_out/launch_trace_c3_metis.txt 2>&1
tail -c 600 gpurun_out/bench_c3_full.json
