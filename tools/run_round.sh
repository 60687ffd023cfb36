set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r17_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r17_bench_c3.json 2> gpurun_out/r17_bench_c3.err
python bench.py --workload c4 --ordering metis --steps 5 --no-cpu-baseline > gpurun_out/r17_bench_c4.json 2> gpurun_out/r17_bench_c4.err
