#!/usr/bin/env python
"""SASS of a source-line range with executed counts (first launch)   python tools/ncu_sass_lines.py rep kernel lo hi"""
import csv
import subprocess
import sys

rep, kern, lo, hi = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hs = [i for i, r in enumerate(rows) if "Instructions Executed" in r]
h = hs[0]
ci = rows[h].index("Instructions Executed")
cur, seen, tot = None, set(), 0
for r in rows[h + 1:hs[1] if len(hs) > 1 else len(rows)]:
    if len(r) <= ci:
        continue
    if r[0] != "":
        try:
            cur = int(r[0])
        except ValueError:
            cur = None
        if cur and lo <= cur <= hi:
            print(f"L{cur}: {r[1].strip()[:120]}")
        continue
    if cur and lo <= cur <= hi and r[ci].isdigit():
        key = (r[2], r[3])
        if key in seen:
            continue
        seen.add(key)
        tot += int(r[ci])
        print(f"    {int(r[ci]):9d}  {r[3].strip()[:100]}")
print("total", tot)
