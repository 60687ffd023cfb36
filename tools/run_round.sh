set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r15_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r15_bench_c3.json 2> gpurun_out/r15_bench_c3.err
python bench.py --workload c4 --ordering metis --steps 5 --no-cpu-baseline > gpurun_out/r15_bench_c4.json 2> gpurun_out/r15_bench_c4.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r15_bench_c3_2gpu.json 2> gpurun_out/r15_bench_c3_2gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --workload c4 > gpurun_out/r15_bench_c4_2gpu.json 2> gpurun_out/r15_bench_c4_2gpu.err
