#!/bin/bash
# tensor-core kNN: parity of the forced MMA path on the matcher tests, then A/B of the two workloads
mkdir -p gpurun_out
MSLAM_KNN_MMA=2 timeout 300 python -m pytest tests -m gpu -q -x --timeout 120 -k "match or tie or knn or frontend or db_" > gpurun_out/pytest_mma.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_mma.log
tail -25 gpurun_out/pytest_mma.log
if grep -q "pytest rc=0" gpurun_out/pytest_mma.log; then
  for m in 0 1; do
    MSLAM_KNN_MMA=$m timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_mma$m.json 2> gpurun_out/bench_mma$m.err
    python - <<PY
import json
d=json.loads(open('gpurun_out/bench_mma$m.json').read().strip().splitlines()[-1])
print('c2 mma=$m', d['value'], d['ms_per_step'], d.get('stage_ms'))
PY
    MSLAM_KNN_MMA=$m timeout 300 python bench.py --workload dist --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dist_mma$m.json 2> gpurun_out/bench_dist_mma$m.err
    python - <<PY
import json
d=json.loads(open('gpurun_out/bench_dist_mma$m.json').read().strip().splitlines()[-1])
print('dist mma=$m', d['value'], d['ms_per_step'], d.get('stage_ms'))
PY
  done
fi
