#!/usr/bin/env python
"""per-source-line instruction / stall-sample totals of one kernel from an .ncu-rep captured with --import-source on
   python tools/ncu_lines.py gpurun_out/prof.ncu-rep k_dist_fast [top]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hs = [i for i, r in enumerate(rows) if "Instructions Executed" in r]
h = hs[0]
end = hs[1] if len(hs) > 1 else len(rows)
hdr = rows[h]
ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
items = []
for r in rows[h + 1:end]:
    if len(r) < len(hdr) or r[0] == "":  # sass rows have an empty line number
        continue
    try:
        items.append((int(r[ci] or 0), int(r[cs] or 0), r[0], r[1].strip()[:110]))
    except ValueError:
        continue
tot = sum(i[0] for i in items) or 1
stot = sum(i[1] for i in items) or 1
print(f"total warp instructions {tot}")
for n, s, l, src in sorted(items, reverse=True)[:top]:
    print(f"{n / tot:6.3f} {s / stot:6.3f}  L{l:>5s}  {src}")
