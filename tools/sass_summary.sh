#!/bin/bash
# SASS of the two step kernels from the built library: instruction mix, then the listing (profiles/r02/sass_*.txt).
# usage: tools/sass_summary.sh <out dir>
O=${1:-profiles/r02}
for k in front_step schur_step; do
  cuobjdump -sass gtsam_petercdev_b200/libgtsam_b200.so | awk -v k="${k}_kernel" '/Function : /{f = index($0, k) > 0} f' > /tmp/sass_$k.txt
  {
    echo "# cuobjdump -sass gtsam_petercdev_b200/libgtsam_b200.so, function gtb::${k}_kernel (sm_100a): instruction mix, then the listing"
    echo "# FP64 tensor instruction = DMMA (mma.sync.m8n8k4.f64); asynchronous copies = LDGSTS (cp.async); hand-over adds = REDG (fire-and-forget);"
    echo "# no UTMALDG / UTCMMA / LDTM: tcgen05 has no f64 kind, TMEM does not apply to this path"
    grep -E '^\s+/\*[0-9a-f]{4,5}\*/' /tmp/sass_$k.txt | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//; s/^@!?U?P[0-9T]+\s+//' | awk '{split($1, a, "."); print a[1] (a[2] ~ /^(8|E|64|128|SYNC|ARV)/ ? "." a[2] : "")}' | sort | uniq -c | sort -rn | awk '{printf "#   %-28s %s\n", $2, $1}'
    cat /tmp/sass_$k.txt
  } > $O/sass_$k.txt
  echo "$k: $(grep -c DMMA /tmp/sass_$k.txt) DMMA, $(grep -c LDGSTS /tmp/sass_$k.txt) LDGSTS, $(grep -c 'REDG\|RED\.' /tmp/sass_$k.txt) RED, $(grep -c 'ATOMG\|ATOM\.' /tmp/sass_$k.txt) ATOM, $(grep -c 'BAR\.' /tmp/sass_$k.txt) BAR"
done
