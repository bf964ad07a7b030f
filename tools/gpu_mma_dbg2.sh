#!/bin/bash
# round 2: the switch-off experiments of tools/gpu_mma_dbg.sh on a -DMSLAM_MMA_DEBUG build (build/libmslam_b200_mmadbg.so), for the
# single-CTA (MSLAM_MMA_CG=1) and the CTA-pair (=2) forms of k_knn2_mma<true>.  Debug modes compute wrong scores by design.
mkdir -p gpurun_out
export MSLAM_BENCH_SKIP_INT_PIPE=1 MSLAM_BENCH_KF_PER_GPU=2500 MSLAM_B200_LIB=build/libmslam_b200_mmadbg.so
for cg in 1 2; do for dbg in ${DBGS:-0 4 1 8 2 5}; do
  MSLAM_MMA_CG=$cg MSLAM_MMA_DEBUG=$dbg timeout 200 python bench.py --workload c5 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/dbg.json 2> gpurun_out/dbg.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/dbg.json').read().strip().splitlines()[-1])
    print('cta_group=$cg debug=$dbg', round(d['value']/1e12,3), 'T pairs/s', round(d['ms_per_step'],2), 'ms')
except Exception as e:
    print('cta_group=$cg debug=$dbg failed', e, open('gpurun_out/dbg.err').read()[-200:])
PY
done; done
