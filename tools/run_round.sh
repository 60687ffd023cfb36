set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r22_pytest.log
for cfg in "c1 metis" "c1 colamd" "c2 metis" "c2 colamd" "c3 metis"; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps 10 --no-cpu-baseline > gpurun_out/r22_bench_$1_$2.json 2>/dev/null
done
