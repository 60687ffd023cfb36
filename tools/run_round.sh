set -x
mkdir -p gpurun_out
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_c3_full.json 2> gpurun_out/bench_c3_full.err
for cfg in "c3 colamd" "c3 schur" "c1 colamd" "c1 metis" "c2 colamd" "c2 metis"; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps 10 > gpurun_out/bench_$1_$2.json 2> gpurun_out/bench_$1_$2.err
done
python bench.py --workload c4 --ordering metis --steps 5 --no-cpu-baseline > gpurun_out/bench_c4_metis.json 2> gpurun_out/bench_c4_metis.err
python bench.py --impl reference --workload c4 --steps 1 --warmup 0 > gpurun_out/bench_c4_ref.json 2> gpurun_out/bench_c4_ref.err
GTB_DEBUG=8 python tools/launch_trace.py c3 metis > gpurun_out/launch_trace_c3_metis.txt 2>&1
tail -2 gpurun_out/smoke.log
