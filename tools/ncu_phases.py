#!/usr/bin/env python
"""k_fast's instructions and stall samples by phase, from the source page of an .ncu-rep captured with --import-source on
   python tools/ncu_phases.py gpurun_out/r2_prof_fast.ncu-rep [launch index]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:k_fast", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hs = [i for i, r in enumerate(rows) if "Instructions Executed" in r and r[0] == "Line No"]
# per launch: one section per source file (the kernel's own file first, then CUDA headers inlined into it)
per = max(1, len(hs) // max(1, sum(1 for i in hs if rows[i + 1][0:1] and rows[i + 1][1].startswith("asm(lop3")) or 1))
h = hs[which * per]
end = hs[which * per + 1] if which * per + 1 < len(hs) else len(rows)
hdr = rows[h]
ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
src = []
for r in rows[h + 1:end]:
    if len(r) < len(hdr) or r[0] == "":
        continue
    try:
        src.append((int(r[0]), r[1].strip(), int(r[ci] or 0), int(r[cs] or 0)))
    except ValueError:
        continue
tot = sum(x[2] for x in src) or 1
stot = sum(x[3] for x in src) or 1
lines = open("modular-slam_b200/csrc/orb_kernels.cu").read().splitlines()


def find(s):
    return next(i + 1 for i, l in enumerate(lines) if s in l)


marks = [("staging", find("which level does this tile belong to")), ("phase A compare+run9", find("---- phase A: classify")),
         ("queue append", find("only positions that can matter")), ("phase B score", find("---- phase B: exact cornerScore")),
         ("phase C nms", find("---- phase C: 3x3 strict NMS")), ("output", find("const int n_out = s_on;")), ("end", find("// K4  select"))]
helpers = [("any_run9", find("__device__ __forceinline__ unsigned any_run9"), find("// the 4 pixels starting dx")),
           ("ring4", find("// the 4 pixels starting dx"), find("// exact cornerScore of TWO pixels")),
           ("corner_score2", find("// exact cornerScore of TWO pixels"), find("__device__ __forceinline__ unsigned fast_smem_u32")),
           ("lop3 helpers", find("__device__ __forceinline__ unsigned lop3_maj_a_nb_c"), find("// bit 7 of each byte of the result")),
           ("mbarrier", find("__device__ __forceinline__ unsigned fast_smem_u32"), find("#ifndef MSLAM_FAST_CTAS"))]
agg, sagg = collections.Counter(), collections.Counter()
for ln, txt, n, s in src:
    name = None
    for nm, a, b in helpers:
        if a <= ln < b:
            name = nm
    if name is None:
        for (nm, a), (_, b) in zip(marks, marks[1:]):
            if a <= ln < b:
                name = nm
    agg[name or "other"] += n
    sagg[name or "other"] += s
print(f"total warp instructions {tot}, samples {stot}")
for k, v in agg.most_common():
    print(f"{k:22s} inst {v / tot:6.3f}  samples {sagg[k] / stot:6.3f}")
