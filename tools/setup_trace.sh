#!/bin/bash
# Where the CLI's set-up time goes, by the driver's hardware-queue count (CUDA_DEVICE_MAX_CONNECTIONS): SSN_TRACE_SETUP lines
# of ssn_create on a small pair, three runs per setting.  Output: gpurun_out/setup_trace.log
set -u
mkdir -p gpurun_out /tmp/st
python - <<'PY'
import sys
sys.path.insert(0, ".")
from somatic_sniper_b200 import cli_tools as ct
ct.make_pair_fast("/tmp/st", [("chr1", 4000000), ("chr2", 4000000)], 60, 30, seed=5, level=1)
PY
CLI=somatic_sniper_b200/bin/bam-somaticsniper
[ -x "$CLI" ] || CLI=$(python -c "from somatic_sniper_b200 import cli_tools as ct; print(ct.CLI)")
{
for conn in 8 16 32; do
  for rep in 1 2 3; do
    echo "== CUDA_DEVICE_MAX_CONNECTIONS=$conn run $rep"
    CUDA_DEVICE_MAX_CONNECTIONS=$conn SSN_TRACE_SETUP=1 SSN_STATS=/tmp/st/stats.json "$CLI" -f /tmp/st/ref.fa /tmp/st/tumor.bam /tmp/st/normal.bam /tmp/st/out.txt 2>&1 | grep -v "^Preparing\|^Normal\|^Tumor\|^Using" | tail -20
    python -c "import json; s=json.load(open('/tmp/st/stats.json')); print('setup_s', s['setup_s'], 'gpu_setup_s', s['gpu_setup_s'], 'wall_s', s['wall_s'], 'col/s', s['columns_per_s'])"
  done
done
} > gpurun_out/setup_trace.log 2>&1
grep -E "^==|setup_s" gpurun_out/setup_trace.log
