#!/bin/bash
# Round-end pass on ONE GPU: GPU tests, smoke, default bench (with the CPU baseline), the other workloads / orderings,
# launch trace, per-level times, ncu launch list with DRAM bytes, ncu --set full of the two contraction kernels.
set -x
mkdir -p gpurun_out/r02
O=gpurun_out/r02
rm -f gpurun_out/parity_deltas.txt
(time timeout 900 python -m pytest tests -m gpu -x -q) > $O/pytest_gpu.log 2>&1
cp gpurun_out/parity_deltas.txt gpurun_out/dropin_test.txt $O/ 2>/dev/null
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1
python bench.py > $O/bench_c3_full.json 2> $O/bench_c3_full.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c3_ref.json 2> $O/bench_c3_ref.err
for cfg in "c1 metis 30" "c2 metis 30" "c2_real metis 30" "c1 colamd 20" "c2 colamd 20" "c2 nested_dissection 20" "c3 colamd 10" "c3 schur 10" "c4 metis 5"; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps $3 > $O/bench_$1_$2.json 2> $O/bench_$1_$2.err
done
# ISAM2 row: the City10000 head (1,264 updates) with the reference on the same host, then 10,000 updates, both arms
python bench.py --workload c5 > $O/bench_c5_isam2_head1264.json 2> $O/bench_c5_head.err
python bench.py --workload c5 --updates 10000 --no-cpu-baseline > $O/bench_c5_isam2_10000.json 2> $O/bench_c5_10000.err
python bench.py --workload c5 --updates 10000 --impl reference > $O/bench_c5_isam2_10000_ref.json 2> $O/bench_c5_10000_ref.err
GTB_DEBUG=8 GTB_TRACE_STEPS=1 python tools/launch_trace.py c3 metis > $O/launch_trace_c3_metis.txt 2>&1
python tools/level_times.py c3 metis > $O/level_times_c3_metis.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv \
  --log-file $O/ncu_launches_c3_metis.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'front_step_kernel|schur_step_kernel' -s 160 -c 24 -o $O/ncu_full_c3 \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_full.log 2>&1
cat $O/pytest_gpu.log | tail -6; tail -2 $O/smoke.log
for f in $O/bench_c*.json; do python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"), d.get("cpu_baseline",{}).get("value"))
except Exception as e: print(sys.argv[1], "ERR", e)
P
done
