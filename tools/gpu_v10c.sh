#!/bin/bash
mkdir -p gpurun_out
for wl in c3 c4; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.log 2> gpurun_out/bench_$wl.err; echo "$wl rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_$wl.log').read().strip().splitlines()[-1])
print('$wl', d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(v['ms_per_step'],4) for k,v in d['roofline']['stages'].items()})
PY
done
NCU_SKIP=0 NCU_COUNT=40 bash tools/gpu_profile.sh > gpurun_out/profile_c2.log 2>&1
tail -3 gpurun_out/profile_c2.log
