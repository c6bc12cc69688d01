#!/usr/bin/env python
"""Aggregate an ncu report's per-SASS metrics to CUDA source lines for one kernel.
usage: ncu_lines.py report.ncu-rep kernel_regex [top_n]"""
import csv
import subprocess
import sys
from collections import defaultdict

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
agg = defaultdict(lambda: [0, 0, ""])
cur_file, hdr, launches = "", None, 0
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif len(r) >= 2 and r[0] == "Function Name":
        launches += 1
    elif len(r) > 8 and r[0] == "Line No":
        hdr = r
        i_line, i_src, i_ex, i_st = 0, 1, hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
    elif hdr and len(r) > 8 and r[0].isdigit():
        try:
            a = agg[(cur_file, int(r[0]))]
            a[0] += int(r[i_ex] or 0)
            a[1] += int(r[i_st] or 0)
            a[2] = r[i_src].strip()[:110]
        except ValueError:
            pass
tot = sum(v[0] for v in agg.values())
tst = sum(v[1] for v in agg.values())
print(f"total warp-instructions {tot}  stall samples {tst}")
for (f, ln), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100 * v[0] / max(tot, 1):5.1f}% inst {100 * v[1] / max(tst, 1):5.1f}% stall  {f}:{ln:<4d} {v[2]}")
