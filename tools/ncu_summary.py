#!/usr/bin/env python
"""Summarise ncu outputs brought back in gpurun_out/ into profiles/ (tracked).
  python tools/ncu_summary.py launches gpurun_out/launches.csv            -> per-kernel count / total / avg / share
  python tools/ncu_summary.py full gpurun_out/prof.ncu-rep                -> one row per captured launch, key metrics
"""
import collections
import csv
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
           "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
           "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "l1tex__m_xbar2l1tex_read_bytes.sum.per_second"]


def launches(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        n = r[ki].split("(")[0]
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += float(r[vi].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    print(f"{'kernel':60s} {'n':>5s} {'total_us':>10s} {'avg_us':>9s} {'share':>6s}")
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n[:60]:60s} {a[0]:5d} {a[1] / 1e3:10.1f} {a[1] / a[0] / 1e3:9.1f} {a[1] / tot:6.3f}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    short = [m.replace(".avg.pct_of_peak_sustained_", "%").replace("sm__inst_executed_pipe_", "pipe_") for m in METRICS]
    print("kernel,grid," + ",".join(short))
    for r in rows[2:]:
        vals = []
        for m in METRICS:
            v = r[hdr.index(m)] if m in hdr else ""
            try:
                v = f"{float(v):.4g}"
            except ValueError:
                pass
            vals.append(v)
        print(r[hdr.index("Kernel Name")].split("(")[0] + "," + r[hdr.index("Grid Size")].replace(",", "x").replace(" ", "")
              + "," + ",".join(vals))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
