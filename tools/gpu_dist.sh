#!/bin/bash
# dist detector: GPU parity tests + stage timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout 300 > gpurun_out/pytest_dist.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist.log
tail -40 gpurun_out/pytest_dist.log
timeout 600 python tools/dist_timing.py > gpurun_out/dist_timing.log 2>&1; cat gpurun_out/dist_timing.log
