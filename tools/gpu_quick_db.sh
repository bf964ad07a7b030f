#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_knn_tensor.py tests/test_sharded_db_gpu.py tests/test_gpu_parity.py -m gpu -x -q -k "db or knn or tensor or sharded or match" 2>&1 | tail -3
for i in 1 2; do
MSLAM_BENCH_SKIP_INT_PIPE=1 timeout 300 python bench.py --workload c5 --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value']/1e12, d['ms_per_step'], d['roofline'].get('frac'))"
done
