#!/bin/bash
# single-frame path (what the reference's frontend calls: one detect + one match per frame): wall times with and without the
# latency path, the per-stage timeline inside the captured graph, the breakdown of a call, and warm per-kernel times
mkdir -p gpurun_out
timeout 200 python tools/latency.py 2>&1 | tail -1
MSLAM_B200_LATENCY_PATH=0 timeout 200 python tools/latency.py 2>&1 | tail -1
MSLAM_B200_LAT_TRACE=2 timeout 200 python tools/latency.py 2>&1 | head -12
timeout 200 python tools/latency_breakdown.py 2>&1 | tail -1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --launch-skip 400 -c 40 --csv --log-file gpurun_out/lat_launches_warm.csv python tools/latency.py > gpurun_out/lat_ncu.log 2>&1
echo "launch list rc=$?"
