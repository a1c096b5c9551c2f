#!/bin/bash
# Whole-genome tiling run (BASELINE.json configs[4]) on a multi-GPU box: diagnostics first, then tools/bench_cli.py.
# usage: tools/run_wg.sh <contigs> <len> <gpu list> <out json>
set -u
C=${1:-32}; L=${2:-8000000}; G=${3:-1,2,4,8}; OUT=${4:-gpurun_out/r2_cli_wg.json}
echo "cores $(nproc)  mem $(free -g | awk 'NR==2{print $2" GB total, "$7" GB available"}')  gpus $(nvidia-smi -L | wc -l)  tmp $(df -h /tmp | awk 'NR==2{print $4}') free"
ulimit -c 0
python tools/bench_cli.py --fast --contigs $C --len $L --gpus $G --reference split --repeat 2 --workdir /tmp/wg --out $OUT > ${OUT%.json}.log 2>&1
rc=$?
echo "bench_cli rc=$rc"
tail -c 1500 ${OUT%.json}.log
if [ $rc -ne 0 ]; then
    dmesg 2>/dev/null | tail -5
    SSN_STATS=/tmp/wg/st.json somatic_sniper_b200/bin/bam-somaticsniper -f /tmp/wg/ref.fa /tmp/wg/tumor.bam /tmp/wg/normal.bam /tmp/wg/direct.out; echo "direct rc=$?"; cat /tmp/wg/st.json 2>/dev/null
    SSN_THREADS=14 SSN_STATS=/tmp/wg/st.json somatic_sniper_b200/bin/bam-somaticsniper -f /tmp/wg/ref.fa /tmp/wg/tumor.bam /tmp/wg/normal.bam /tmp/wg/direct.out; echo "direct (14 threads) rc=$?"; cat /tmp/wg/st.json 2>/dev/null
fi
