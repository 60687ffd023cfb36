#!/bin/bash
# The bench lines only (no tests, no ncu), one GPU: refreshes gpurun_out/r02/bench_*.json after a host-side change.
set -x
O=gpurun_out/r02
mkdir -p $O
python bench.py > $O/bench_c3_full.json 2> $O/bench_c3_full.err
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_c3_ref.json 2> $O/bench_c3_ref.err
for cfg in "c1 metis 30" "c2 metis 30" "c2_real metis 30" "c1 colamd 20" "c2 colamd 20" "c2 nested_dissection 20" "c3 colamd 10" "c3 schur 10" "c4 metis 5"; do
  set -- $cfg
  python bench.py --workload $1 --ordering $2 --steps $3 > $O/bench_$1_$2.json 2> $O/bench_$1_$2.err
done
python tools/marginals_bench.py c1 > $O/marginals_c1.json 2> $O/marginals_c1.err
python tools/marginals_bench.py c2 > $O/marginals_c2.json 2> $O/marginals_c2.err
for f in $O/bench_c[1234]*.json; do python - "$f" <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d["n_gpus"], round(d["value"],2), round(d["ms_per_step"],3), round(d.get("e2e",{}).get("value",0),2), d.get("cpu_baseline",{}).get("value"), d.get("roofline",{}).get("traffic"))
except Exception as e: print(sys.argv[1], "ERR", e)
P
done
cat $O/marginals_c1.json $O/marginals_c2.json | cut -c1-300
