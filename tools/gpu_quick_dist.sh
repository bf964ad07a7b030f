#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist_gpu.py tests/test_reference_build.py -m gpu -x -q 2>&1 | tail -3
for i in 1 2; do
timeout 300 python bench.py --workload dist --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], round(d.get('e2e',{}).get('value',0)), {k:round(v['ms_per_step'],4) for k,v in d['roofline']['stages'].items()})"
done
