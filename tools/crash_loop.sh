#!/bin/bash
# run the CLI repeatedly on a pair until it fails; print the crash trace
D=$1; N=${2:-12}
for i in $(seq 1 $N); do
  SSN_STATS=$D/st.json somatic_sniper_b200/bin/bam-somaticsniper -f $D/ref.fa $D/tumor.bam $D/normal.bam $D/loop.out 2> $D/loop.err; rc=$?
  echo "run $i rc=$rc $(python -c "import json;d=json.load(open('$D/st.json'));print(d['wall_s'],d['columns_per_s'])" 2>/dev/null)"
  if [ $rc -ne 0 ]; then cat $D/loop.err | tail -40; grep -m3 "libssn" /proc/self/maps; break; fi
done
