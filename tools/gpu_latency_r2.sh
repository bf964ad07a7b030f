#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/latency.py 2>&1 | tail -1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --workload dist --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d.get('e2e',{}).get('value'))"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --launch-skip 100 -c 16 --csv --log-file gpurun_out/lat_dist_warm.csv python tools/lat_dist.py > gpurun_out/lat_ncu.log 2>&1
grep k_dist_tree gpurun_out/lat_dist_warm.csv | head -2 | cut -c1-40,230-
