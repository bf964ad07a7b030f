#!/bin/bash
mkdir -p gpurun_out
for m in 0 1; do
  for wl in c2 dist c5; do
    MSLAM_KNN_MMA=$m timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${wl}_mma$m.json 2> gpurun_out/bench_${wl}_mma$m.err
    python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_${wl}_mma$m.json').read().strip().splitlines()[-1])
    print('$wl mma=$m', d['value'], d['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in d['roofline'].get('stages',{}).items()})
except Exception as e:
    print('$wl mma=$m failed', e)
PY
  done
done
