#!/usr/bin/env python
"""k_fast's instructions and stall samples by phase, from the source page of an .ncu-rep captured with --import-source on
   python tools/ncu_phases.py gpurun_out/r2_prof_fast.ncu-rep [launch index]
The report has one section per (launch, source file): the kernel's own file (orb_kernels.cu: phases by line markers), the
shared helpers (fast_simd.cuh: by function) and CUDA headers inlined into it."""
import collections
import csv
import os
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:k_fast", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hs = [i for i, r in enumerate(rows) if "Instructions Executed" in r and r[0] == "Line No"]
files = []
for h in hs:
    f = next((r[1] for r in rows[max(0, h - 4):h] if r and r[0] == "File Path"), "?")
    files.append(os.path.basename(f))
# a launch = the run of sections up to the next repetition of the first file
first = files[0]
starts = [i for i, f in enumerate(files) if f == first]
lo = starts[which]
hi = starts[which + 1] if which + 1 < len(starts) else len(hs)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def marks_of(path, marks):
    lines = open(os.path.join(ROOT, "modular-slam_b200", "csrc", path)).read().splitlines()
    out = []
    for name, text in marks:
        out.append((name, next(i + 1 for i, l in enumerate(lines) if text in l)))
    return out


PHASES = {
    "orb_kernels.cu": marks_of("orb_kernels.cu", [("staging", "which level does this tile belong to"), ("phase A compare", "---- phase A: classify"),
                                                  ("queue append", "only positions that can matter"),
                                                  ("phase B decode", "---- phase B: exact cornerScore"), ("phase C nms", "---- phase C: 3x3 strict NMS"),
                                                  ("output", "const int n_out = s_on;"), ("end", "// K4  select")]),
    "fast_simd.cuh": marks_of("fast_simd.cuh", [("lop3 helpers", "lop3_maj_a_nb_c(unsigned a"), ("any_run9", "__device__ __forceinline__ unsigned any_run9"),
                                                ("ring4", "__device__ __forceinline__ unsigned ring4"), ("corner_score2", "template <int S"),
                                                ("end", "} // namespace mslam_b200")]),
}
agg, sagg = collections.Counter(), collections.Counter()
for si in range(lo, hi):
    h = hs[si]
    end = hs[si + 1] if si + 1 < len(hs) else len(rows)
    hdr = rows[h]
    ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
    marks = PHASES.get(files[si])
    for r in rows[h + 1:end]:
        if len(r) < len(hdr) or r[0] == "":
            continue
        try:
            ln, n, s = int(r[0]), int(r[ci] or 0), int(r[cs] or 0)
        except ValueError:
            continue
        name = files[si]
        if marks:
            name = "(before the kernel)"
            for (nm, a), (_, b) in zip(marks, marks[1:]):
                if a <= ln < b:
                    name = nm
        agg[name] += n
        sagg[name] += s
tot = sum(agg.values()) or 1
stot = sum(sagg.values()) or 1
print(f"launch {which}: total warp instructions {tot}, samples {stot}")
for k, v in agg.most_common():
    print(f"{k:28s} inst {v / tot:6.3f}  samples {sagg[k] / stot:6.3f}")
