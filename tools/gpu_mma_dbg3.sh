#!/bin/bash
# late round 2: switch-off experiments on the N = 256 database-scoring kernel (k_knn2_mma_wide<true, 8>), -DMSLAM_MMA_DEBUG build
# (build/libmslam_b200_mmadbg.so): 0 = as shipped, 4 = epilogue only hands the accumulators back (MMAs + feed alone),
# 1 = one K step per tile (epilogue alone), 5 = both (feed + hand-overs alone).  Debug modes compute wrong scores by design.
mkdir -p gpurun_out
export MSLAM_BENCH_SKIP_INT_PIPE=1 MSLAM_BENCH_KF_PER_GPU=2500 MSLAM_B200_LIB=build/libmslam_b200_mmadbg.so
for dbg in ${DBGS:-0 4 1 5}; do
  MSLAM_MMA_DEBUG=$dbg timeout 200 python bench.py --workload c5 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/dbg.json 2> gpurun_out/dbg.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/dbg.json').read().strip().splitlines()[-1])
    print('debug=$dbg', round(d['value']/1e12,3), 'T pairs/s', round(d['ms_per_step'],2), 'ms')
except Exception as e:
    print('debug=$dbg failed', e, open('gpurun_out/dbg.err').read()[-200:])
PY
done
