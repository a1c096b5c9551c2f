#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` dump (SASS view) by CUDA source line / source region.

usage: sass_by_line.py <source.csv> <nvdisasm -g -c output> <kernel mangled-name substring> [regions.txt]
The disassembly of the same build gives offset -> (file, line); ncu gives per-instruction executed counts
and stall samples in the same order.  Prints per-region and top per-line totals.
"""
import csv, re, sys
from collections import defaultdict

src_csv, dis, kname = sys.argv[1:4]
# --- disassembly: offset -> line
lines = {}
cur = None
inside = False
for ln in open(dis, errors="replace"):
    if ln.startswith(".text."):
        inside = kname in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        if "inlined at" not in ln or cur is None:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        lines[int(m.group(1), 16)] = (cur, m.group(2).strip())
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ia, ii, isamp, isrc = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
body = rows[2:]
base = int(body[0][ia], 16)
per_line = defaultdict(lambda: [0, 0])
tot_i = tot_s = 0
unmapped = 0
for r in body:
    off = int(r[ia], 16) - base
    n, s = int(r[ii] or 0), int(r[isamp] or 0)
    tot_i += n; tot_s += s
    key = lines.get(off, (None, ""))[0]
    if key is None:
        unmapped += n
        key = ("?", 0)
    per_line[key][0] += n; per_line[key][1] += s
print("total warp instructions %d, samples %d, unmapped %d" % (tot_i, tot_s, unmapped))
regions = []
if len(sys.argv) > 4:
    for ln in open(sys.argv[4]):
        f = ln.split()
        if len(f) >= 4:
            regions.append((f[0], int(f[1]), int(f[2]), " ".join(f[3:])))
if regions:
    agg = defaultdict(lambda: [0, 0])
    for (f, l), (n, s) in per_line.items():
        name = "other"
        for rf, a, b, nm in regions:
            if rf == f and a <= l <= b:
                name = nm
                break
        agg[name][0] += n; agg[name][1] += s
    print("%-34s %12s %7s %9s %7s" % ("region", "warp-inst", "%", "samples", "%"))
    for nm, (n, s) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print("%-34s %12d %6.1f%% %9d %6.1f%%" % (nm, n, 100.0 * n / tot_i, s, 100.0 * s / max(tot_s, 1)))
print("top lines:")
for (f, l), (n, s) in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:40]:
    print("  %s:%d  inst %d (%.1f%%)  samples %d (%.1f%%)" % (f, l, n, 100.0 * n / tot_i, s, 100.0 * s / max(tot_s, 1)))
