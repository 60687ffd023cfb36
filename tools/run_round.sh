set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r4_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r4_bench_c3.json 2> gpurun_out/r4_bench_c3.err
python bench.py --workload c4 --ordering metis --steps 5 --no-cpu-baseline > gpurun_out/r4_bench_c4.json 2> gpurun_out/r4_bench_c4.err
tail -5 gpurun_out/r4_pytest.log; tail -c 300 gpurun_out/r4_bench_c3.err
