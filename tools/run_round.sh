set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r16_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r16_bench_c3.json 2> gpurun_out/r16_bench_c3.err
