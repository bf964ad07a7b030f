#!/bin/bash
# per-kernel durations of single-frame calls (launch list only; numbers under ncu are never bench values)
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/lat_launches.csv python tools/latency.py > gpurun_out/lat_ncu.log 2>&1
echo "rc=$?"
tail -3 gpurun_out/lat_ncu.log
