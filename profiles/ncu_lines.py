#!/usr/bin/env python
"""Per-source-line instruction counts / stall samples of an ncu report (needs -lineinfo + --import-source on):
python profiles/ncu_lines.py x.ncu-rep [top=40]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
hdr, cur, out, seen_kernel = None, None, [], 0
for r in csv.reader(io.StringIO(txt)):
    if not r:
        continue
    if r[0] == "Function Name":
        seen_kernel += 1
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) > 6 and r[0].isdigit():
        def num(x):
            return int(x) if x.strip().isdigit() else 0
        out.append((num(r[hdr.index("Instructions Executed")]), num(r[hdr.index("# Samples")]), cur, r[0], r[1].strip()[:100]))
out.sort(reverse=True)
tot = sum(o[0] for o in out)
print("total warp instructions over %d function blocks: %d" % (seen_kernel, tot))
for o in out[:top]:
    print("%9d %5.1f%% samples=%4d %s:%s  %s" % (o[0], 100.0 * o[0] / max(tot, 1), o[1], o[2], o[3], o[4]))
