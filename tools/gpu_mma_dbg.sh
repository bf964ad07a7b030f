#!/bin/bash
# timing experiments on the tensor-core DB scoring kernel (results of debug modes are wrong by design; only ms matter)
mkdir -p gpurun_out
export MSLAM_BENCH_SKIP_INT_PIPE=1 MSLAM_BENCH_KF_PER_GPU=2500
for dbg in ${DBGS:-0 1 2 3 4 5}; do
  MSLAM_MMA_DEBUG=$dbg timeout 300 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/dbg_$dbg.json 2> gpurun_out/dbg_$dbg.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/dbg_$dbg.json').read().strip().splitlines()[-1])
    print('debug=$dbg', round(d['value']/1e12,3), 'T pairs/s', round(d['ms_per_step'],2), 'ms')
except Exception as e:
    print('debug=$dbg failed', e, open('gpurun_out/dbg_$dbg.err').read()[-300:])
PY
done
