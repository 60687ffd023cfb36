set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r18_pytest.log
for cfg in "c1 metis" "c1 colamd" "c2 metis" "c2 colamd"; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps 10 --no-cpu-baseline > gpurun_out/r18_bench_$1_$2.json 2>/dev/null
  GTB_THIN_THREADS=128 GTB_SMALL_MUL=1 python bench.py --workload $1 --ordering $2 --steps 10 --no-cpu-baseline > gpurun_out/r18_bench_$1_$2_old.json 2>/dev/null
done
