# usage: run_multi.sh N [c4]   (on a box with N GPUs): the tree-partitioned bench for C3 (and C4 when asked), as the driver
# launches it; also the N = 1 line of the same box for the ratio
set -x
N=$1
O=gpurun_out/scale
mkdir -p $O
[ "$N" = "2" ] && python bench.py --no-cpu-baseline --steps 10 --warmup 3 > $O/bench_c3_1gpu_on_${N}gpu_box.json 2> $O/bench_c3_1gpu_on_${N}gpu_box.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_c3_${N}gpu.json 2> $O/bench_c3_${N}gpu.err
if [ "$2" = "c4" ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 5 --warmup 3 --workload c4 --no-cpu-baseline > $O/bench_c4_${N}gpu.json 2> $O/bench_c4_${N}gpu.err
fi
tail -c 300 $O/bench_c3_${N}gpu.err
for f in $O/bench_c*_${N}gpu*.json; do python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d["n_gpus"], d["value"], d["ms_per_step"], d.get("e2e",{}).get("value"), d["roofline"]["all_classes_ms_per_step"].get("collectives"))
except Exception as e: print(sys.argv[1], "ERR", e)
P
done
