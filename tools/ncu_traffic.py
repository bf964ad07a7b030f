#!/usr/bin/env python
"""profiles/ncu_traffic.json from an ncu CSV of dram__bytes_read.sum / dram__bytes_write.sum per launch
(tools/gpu_final_r2.sh: `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 90 --csv` on
`bench.py --serial --steps 1 --warmup 3 --no-sub --no-cpu-baseline`): DRAM bytes of ONE c2 step, per stage.  A step is the
launches from one k_gray16 up to the next; the last complete step of the capture is taken (warm kernels, cold L2: ncu flushes
caches between replays).   python tools/ncu_traffic.py gpurun_out/r2_traffic.csv <git rev> > profiles/ncu_traffic.json"""
import collections
import csv
import json
import sys

STAGE = [("k_gray", "gray"), ("k_resize", "resize"), ("k_fast", "fast"), ("k_select", "select"), ("k_blur", "blur"),
         ("k_orient_describe", "describe"), ("k_expand_pm1", "knn"), ("k_knn2", "knn"), ("k_ratio_compact", "compact"),
         ("k_mutual_compact", "compact")]

rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
ki, mi, vi, ui, ii = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit"), hdr.index("ID")
launch = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(r[ui], 1)
    e = launch.setdefault(int(r[ii]), {"name": r[ki].split("(")[0], "bytes": 0.0})
    if "dram__bytes" in r[mi]:
        e["bytes"] += v * scale
seq = list(launch.values())
starts = [i for i, e in enumerate(seq) if "k_gray" in e["name"]]
assert len(starts) >= 2, "need at least one complete step"
a, b = starts[-2], starts[-1]
per = collections.OrderedDict()
names = collections.OrderedDict()
for e in seq[a:b]:
    for pat, st in STAGE:
        if pat in e["name"]:
            per[st] = per.get(st, 0) + e["bytes"]
            names.setdefault(st, []).append(e["name"].split("::")[-1])
            break
print(json.dumps({"_comment": "dram__bytes_read.sum + dram__bytes_write.sum of ONE c2 step (64 frames of 640x480; all launches of a stage), ncu "
                              "--metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none on `bench.py --serial --steps 1 --warmup 3 "
                              "--no-sub --no-cpu-baseline` (tools/gpu_final_r2.sh, tools/ncu_traffic.py); cold L2 per launch: ncu flushes caches "
                              "between replays, so re-reads that the 126 MB L2 serves in the timed pipeline count as DRAM here",
                  "captured_at": sys.argv[2] if len(sys.argv) > 2 else None, "workload": "c2",
                  "launches_in_step": {k: v for k, v in names.items()},
                  "bytes_per_step": {k: int(v) for k, v in per.items()}}, indent=1))
