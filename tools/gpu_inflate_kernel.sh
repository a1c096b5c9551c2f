#!/bin/bash
# GPU box pass over the inflate kernel alone: parity test against zlib, throughput as a reader sees it (H2D + kernel + D2H
# per batch) at several reader counts, and one full ncu capture of the kernel.  Results under gpurun_out/.
# Usage: tools/gpu_inflate_kernel.sh [tag] [ncu: 1|0]
set -u
TAG=${1:-v}
NCU=${2:-1}
mkdir -p gpurun_out
python -m pytest tests/test_inflate.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
python tools/inflate_bench.py --level 1 --batch 64 --threads 1,16,48 > "gpurun_out/inflate_bench_l1_b64_$TAG.json"
python tools/inflate_bench.py --level 1 --batch 96 --threads 1,16,32 > "gpurun_out/inflate_bench_l1_b96_$TAG.json"
python tools/inflate_bench.py --level 6 --batch 128 --threads 16,48 > "gpurun_out/inflate_bench_l6_b128_$TAG.json"
cat gpurun_out/inflate_bench_l*_"$TAG".json | python -c '
import sys, json
for l in sys.stdin:
    r = json.loads(l)
    print("level", r["level"], "batch", r["batch"], [(x["threads"], round(x["GBs"], 2)) for x in r["runs"]])
'
if [ "$NCU" = 1 ]; then
    timeout 300 ncu --set full --clock-control none --import-source on -k regex:ssn_inflate_kernel -s 2 -c 1 -f -o "gpurun_out/inflate_$TAG" \
        python tools/inflate_bench.py --len 2000000 --level 1 --batch 96 --threads 1 --seconds 0.1 > gpurun_out/ncu_inflate.log 2>&1
    tail -2 gpurun_out/ncu_inflate.log
fi
