#!/bin/bash
# A/B of the host pipeline's chunk count:  VAR=MSLAM_B200_DIST_CHUNKS WL=dist bash tools/gpu_ab_chunks.sh
mkdir -p gpurun_out
for N in ${COUNTS:-1 2 4}; do
  env ${VAR:-MSLAM_B200_CHUNKS}=$N python bench.py --workload ${WL:-c2} --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab.log 2> gpurun_out/ab.err
  python -c "
import json;d=json.loads(open('gpurun_out/ab.log').readline());e=d['e2e'];print('${VAR:-MSLAM_B200_CHUNKS}=$N', 'pipelined %.0f (%.3f ms)  sync %.0f (%.3f ms)'%(e['value'],e['ms_per_step'],e['synchronous']['value'],e['synchronous']['ms_per_step']))"
done
