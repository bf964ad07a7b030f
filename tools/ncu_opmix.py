#!/usr/bin/env python
"""opcode mix (executed warp instructions) of one kernel launch from an .ncu-rep with SASS
   python tools/ncu_opmix.py rep.ncu-rep kernel_regex [launch index]"""
import collections
import csv
import re
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hs = [i for i, r in enumerate(rows) if "Instructions Executed" in r]
h = hs[which]
end = hs[which + 1] if which + 1 < len(hs) else len(rows)
ci = rows[h].index("Instructions Executed")
op = collections.Counter()
for r in rows[h + 1:end]:
    if len(r) > ci and r[ci].isdigit():
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[1].strip())
        if m:
            op[m.group(2).split(".")[0]] += int(r[ci])
tot = sum(op.values())
print(f"total warp instructions {tot}")
for k, v in op.most_common(24):
    print(f"{k:10s} {v:12d} {v / tot:6.3f}")
