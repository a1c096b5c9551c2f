#!/usr/bin/env python
"""Prints the SASS (nvdisasm -g -c output) attributed to a source line range of one kernel.
usage: sass_of_lines.py <dis.txt> <kernel substring> <file> <first> <last>"""
import re, sys
dis, kname, fname, a, b = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
inside = False; cur = None; n = 0
for ln in open(dis, errors="replace"):
    if ln.startswith(".text."):
        inside = kname in ln; continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        if "inlined at" not in ln or cur is None:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m and cur and cur[0] == fname and a <= cur[1] <= b:
        print("%s:%d  %s  %s" % (cur[0], cur[1], m.group(1), m.group(2).strip())); n += 1
    elif re.match(r"\s*\.L_x_\d+:", ln) and cur and cur[0] == fname and a <= cur[1] <= b:
        print(ln.strip())
print("instructions:", n)
