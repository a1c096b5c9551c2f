#!/bin/bash
# One GPU box pass over the CLI with the readers inflating on the GPU: the CLI parity tests (default = GPU inflate), then
# the 64 Mb pair with the host decoder and with the GPU decoder on the same files, outputs compared.
# Usage: tools/gpu_inflate_cli.sh [contigs] [gpu list] [variants] [tests: 1|0]     (results under gpurun_out/)
set -u
C=${1:-8}
G=${2:-1}
V=${3:-SSN_GPU_INFLATE=0;SSN_GPU_INFLATE=1}
T=${4:-1}
mkdir -p gpurun_out
if [ "$T" = 1 ]; then
    timeout 900 python -m pytest tests/test_cli_gpu.py tests/test_inflate.py -x -q -m gpu > gpurun_out/cli_tests_gpu_inflate.log 2>&1
    echo "rc=$?" >> gpurun_out/cli_tests_gpu_inflate.log
    tail -3 gpurun_out/cli_tests_gpu_inflate.log
fi
timeout 1500 python tools/bench_cli.py --fast --contigs "$C" --len 8000000 --gpus "$G" --reference none --repeat ${REPEAT:-3} \
    --variants "$V" --out "gpurun_out/cli_inflate_${C}x8mb_gpus_${G//,/_}.json" > /dev/null 2> gpurun_out/cli_inflate.err
echo "bench rc=$?"
tail -5 gpurun_out/cli_inflate.err
python - "gpurun_out/cli_inflate_${C}x8mb_gpus_${G//,/_}.json" <<'EOF'
import json, sys
r = json.load(open(sys.argv[1]))
for k, v in r["runs"].items():
    print(k, "stream %.3g col/s" % v["columns_per_s_stream"], "wall %.2f" % v["wall_s"], "setup %.2f" % v["setup_s"],
          "identical", v.get("identical_to_first_run"), "blocks", v["stats"].get("gpu_inflate_blocks"), "refused", v["stats"].get("gpu_inflate_refused"),
          "all", ["%.3g" % x for x in v["all_stream_columns_per_s"]])
EOF
