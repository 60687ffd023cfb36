set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_fullsize_gpu.py -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r20_pytest.log
