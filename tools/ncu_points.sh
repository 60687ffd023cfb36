set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:point_ -c 10 -f -o gpurun_out/r2_points python bench.py --workload c4 --ordering metis --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2_ncu_points.log 2>&1
ls -la gpurun_out/
