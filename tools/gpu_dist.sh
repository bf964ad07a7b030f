#!/bin/bash
# dist detector: GPU parity tests + plugin test + stage timing + bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dist_gpu.py tests/test_plugin.py -m gpu -q --timeout 300 > gpurun_out/pytest_dist.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist.log
tail -15 gpurun_out/pytest_dist.log
timeout 600 python tools/dist_timing.py > gpurun_out/dist_timing.log 2>&1; cat gpurun_out/dist_timing.log
timeout 600 python bench.py --workload dist --steps 20 --warmup 5 ${BENCH_EXTRA} > gpurun_out/bench_dist.log 2> gpurun_out/bench_dist.err; echo "bench rc=$?" >> gpurun_out/bench_dist.err
python -c "
import json;d=json.loads(open('gpurun_out/bench_dist.log').readline());print(d['value'],d['ms_per_step'],d['e2e']);print({k:round(v['ms_per_step'],4) for k,v in d['roofline']['stages'].items()});print(d['cpu_baseline'])"; tail -5 gpurun_out/bench_dist.err
