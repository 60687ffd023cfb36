#!/bin/bash
# Round bench matrix on one B200: default bench (C3/METIS, with CPU baseline), other workloads/orderings, reference arm,
# ncu launch list + DRAM traffic of the default bench command, --set full captures of the dominant kernels, per-launch
# trace.  Outputs land in gpurun_out/ (copy what should be judged into profiles/rNN/).
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_c3_full.json 2> gpurun_out/bench_c3_full.err
for cfg in "c3 colamd" "c3 schur" "c1 colamd" "c1 metis" "c2 colamd" "c2 metis"; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps 10 > gpurun_out/bench_$1_$2.json 2> gpurun_out/bench_$1_$2.err
done
python bench.py --workload c4 --ordering metis --steps 5 --no-cpu-baseline > gpurun_out/bench_c4_metis.json 2> gpurun_out/bench_c4_metis.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c3_ref.json 2> gpurun_out/bench_c3_ref.err
GTB_DEBUG=8 python tools/launch_trace.py c3 metis > gpurun_out/launch_trace_c3_metis.txt 2>&1
python tools/marginals_bench.py c1 > gpurun_out/marginals_c1.json 2> gpurun_out/marginals_c1.err
python tools/marginals_bench.py c2 > gpurun_out/marginals_c2.json 2> gpurun_out/marginals_c2.err
# ncu: launch list with DRAM bytes (the same command as the bench, fewer steps), then full captures of the top kernels
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv \
  --log-file gpurun_out/ncu_launches_c3_metis.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
# one launch of every top kernel from the middle of a warm step; the report is turned into csv pages here (it is too big to
# travel) and removed
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'front_step|schur_step|point_|backsolve_kernel' -s 330 -c 24 -f \
  -o gpurun_out/ncu_full_c3 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/ncu_full_c3.ncu-rep --page raw --csv > gpurun_out/ncu_full_c3_raw.csv 2>/dev/null
ncu -i gpurun_out/ncu_full_c3.ncu-rep --page source --csv -k regex:front_step -c 1 2>/dev/null | gzip > gpurun_out/ncu_full_c3_front_step_source.csv.gz
rm -f gpurun_out/ncu_full_c3.ncu-rep
gzip -f gpurun_out/ncu_launches_c3_metis.csv
tail -c 600 gpurun_out/bench_c3_full.json
