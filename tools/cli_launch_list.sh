#!/bin/bash
# Launch list of the whole program (ncu, durations only): which kernels bam-somaticsniper runs on a small pair with the
# readers inflating on the GPU, and each kernel's share of the GPU time.  Per-launch times under ncu are serialised and
# cold; the shares are what this is for.  Output: gpurun_out/cli_launches.csv, gpurun_out/cli_launches_summary.txt
set -u
mkdir -p gpurun_out /tmp/lp
python - <<'PY'
import sys
sys.path.insert(0, ".")
from somatic_sniper_b200 import cli_tools as ct
ct.make_pair_fast("/tmp/lp", [("chr1", 3000000), ("chr2", 3000000)], 60, 30, seed=5, level=1)
print(ct.CLI)
PY
CLI=$(python -c "from somatic_sniper_b200 import cli_tools as ct; print(ct.CLI)")
SSN_GPU_INFLATE=1 timeout 170 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/cli_launches.csv \
    "$CLI" -f /tmp/lp/ref.fa /tmp/lp/tumor.bam /tmp/lp/normal.bam /tmp/lp/out.txt > gpurun_out/cli_launches.log 2>&1
echo "ncu rc=$?"
python - <<'PY' | tee gpurun_out/cli_launches_summary.txt
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/cli_launches.csv")) if len(r) > 10]
h = rows[0]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
t, n = collections.Counter(), collections.Counter()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
    k = r[ki].split("(")[0]
    t[k] += v; n[k] += 1
tot = sum(t.values())
print("kernel, launches, total us, share")
for k, v in t.most_common():
    print("%s, %d, %.0f, %.1f%%" % (k, n[k], v, 100 * v / tot))
print("all, %d, %.0f" % (sum(n.values()), tot))
PY
