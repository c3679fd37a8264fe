#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel from an ncu report (needs -lineinfo + --import-source on):

    python tools/ncu_hotlines.py gpurun_out/prof.ncu-rep [units] [top] [kernel regex]

prints warp instructions and stall samples per source line, divided by `units` (e.g. 32768 row-warps)."""
import csv
import subprocess
import sys
import collections

rep = sys.argv[1]
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 14
kern = ["--kernel-name", "regex:" + sys.argv[4]] if len(sys.argv) > 4 else []
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"] + kern, capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
files = collections.OrderedDict()
cur = None
seen_kernel = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1]
        continue
    if r[0] == "Function Name":
        if seen_kernel is None:
            seen_kernel = r[1]
        take = r[1] == seen_kernel
        continue
    if r[0] == "Line No" or cur is None or not take:
        continue
    if r[0].isdigit():  # a source line row: aggregated over its SASS
        try:
            inst = float(r[7])
            samples = float(r[6]) if r[6] not in ("-", "") else 0.0
        except (ValueError, IndexError):
            continue
        d = files.setdefault(cur, {})
        key = (int(r[0]), r[1].strip())
        a = d.get(key, [0.0, 0.0])
        a[0] += inst
        a[1] += samples
        d[key] = a
print("kernel:", seen_kernel)
tot = 0.0
for f, d in files.items():
    s = sum(v[0] for v in d.values())
    tot += s
    print("== %s warp-inst %d per unit %.1f" % (f.split("/")[-1], s, s / units))
    for (ln, src), v in sorted(d.items(), key=lambda kv: -kv[1][0])[:top]:
        print("   %5d %8.1f  samples %6d  %s" % (ln, v[0] / units, v[1], src[:100]))
print("TOTAL per unit", tot / units)
