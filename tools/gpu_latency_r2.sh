#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/latency.py 2>&1 | tail -1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/q_c2.json 2> gpurun_out/q_c2.err; tail -2 gpurun_out/q_c2.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/q_c2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(v['ms_per_step'],4) for k,v in d['roofline']['stages'].items()})
print(d['e2e'].get('plugin_single_frame')); print(d['e2e'].get('plugin_batched'))
P
