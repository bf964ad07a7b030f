#!/usr/bin/env python
"""SASS evidence of the tcgen05 / TMA / mbarrier code in libmslam_b200.so: per kernel, how often each Blackwell mnemonic occurs and
its first occurrence.   python tools/sass_excerpt.py > profiles/<round>_sass_excerpt.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "modular-slam_b200", "lib", "libmslam_b200.so")
WANT = ("UTCIMMA", "UTCBAR", "LDTM", "UBLKCP", "UTMALDG", "SYNCS", "VIMNMX3", "UCGABAR", "UTCATOMSWS", "MEMBAR.ALL.CTA", "FENCE.VIEW.ASYNC")
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
kern, stats = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)", "<anon>").split("(")[0]
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
    if not m or kern is None:
        continue
    ins = m.group(1).strip()
    op = ins.split()[1] if ins.startswith("@") else ins.split()[0]
    for w in WANT:
        if op.startswith(w):
            e = stats.setdefault(kern, collections.OrderedDict()).setdefault(op, [0, ins])
            e[0] += 1
print(f"cuobjdump -sass {os.path.relpath(LIB, ROOT)}  (sm_100a): Blackwell-specific mnemonics per kernel -- count, first occurrence")
for k, ops in stats.items():
    if not any(o.startswith(("UTCIMMA", "UTMALDG", "UBLKCP", "LDTM", "UCGABAR")) for o in ops):
        continue
    print(f"\n{k}")
    for op, (n, ins) in sorted(ops.items()):
        print(f"  {n:4d} x {op:34s} e.g.  {ins[:110]}")
