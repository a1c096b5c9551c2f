#!/bin/bash
# Round-end pass on one GPU box: the whole GPU suite, the bench line, the inflate kernel's throughput (default engine and
# with a stream per request) and the CLI on the 64 Mb pair with host and GPU inflate.  Results under gpurun_out/.
set -u
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/gpu_suite_final.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/gpu_suite_final.log
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_final.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_final.json 2>> gpurun_out/bench_final.err; echo "ref rc=$?"
python tools/inflate_bench.py --level 1 --batch 96 --threads 1,16,32 > gpurun_out/inflate_bench_l1_b96_final.json
SSN_INFLATE_WAVES=32 SSN_INFLATE_WAVE_BLOCKS=96 CUDA_DEVICE_MAX_CONNECTIONS=32 python tools/inflate_bench.py --level 1 --batch 96 --threads 16,32 > gpurun_out/inflate_bench_l1_b96_final_stream_per_request.json
SSN_INFLATE_WAVES=32 SSN_INFLATE_WAVE_BLOCKS=128 CUDA_DEVICE_MAX_CONNECTIONS=32 python tools/inflate_bench.py --level 6 --batch 128 --threads 16,32 > gpurun_out/inflate_bench_l6_b128_final_stream_per_request.json
for f in gpurun_out/inflate_bench_*final*.json; do python -c '
import sys, json
r = json.load(open(sys.argv[1])); print(sys.argv[1], "level", r["level"], "batch", r["batch"], [(x["threads"], round(x["GBs"], 2)) for x in r["runs"]])' "$f"; done
REPEAT=3 bash tools/gpu_inflate_cli.sh 8 1 "SSN_GPU_INFLATE=0;SSN_GPU_INFLATE=1" 0
cp gpurun_out/cli_inflate_8x8mb_gpus_1.json gpurun_out/cli_64mb_host_vs_gpu_inflate_final.json
