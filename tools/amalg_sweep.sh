#!/bin/bash
# Amalgamation sweep: step time per workload / ordering for several padded-area ratios and front-size caps.
run() { # workload ordering ratio maxf
  GTB_AMALGAMATE=$3 GTB_AMALG_MAXF=$4 python bench.py --workload $1 --ordering $2 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); c=d['roofline']['all_classes_ms_per_step']; print('$1 $2 ratio=$3 maxF=$4: %.3f ms/step  levels %d  err %.10g ' % (d['ms_per_step'], d['details']['levels_of_the_numeric_plan'], d['trial_error']), {k:v for k,v in c.items() if v>0.15})"
}
for cfg in "0 512" "1.35 512" "1.35 768" "1.6 768" "1.6 1024" "2.0 1024"; do set -- $cfg; run c3 metis $1 $2; done
for cfg in "0 512" "1.35 512" "1.6 768" "2.0 1024"; do set -- $cfg; run c3 colamd $1 $2; run c3 schur $1 $2; done
for cfg in "0 512" "1.35 512" "1.6 768"; do set -- $cfg; run c4 metis $1 $2; run c2 metis $1 $2; run c2 colamd $1 $2; run c1 colamd $1 $2; done
