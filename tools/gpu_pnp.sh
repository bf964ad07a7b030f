#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_pnp.py -m gpu -q --timeout 300 > gpurun_out/pytest_pnp.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_pnp.log
tail -40 gpurun_out/pytest_pnp.log
timeout 300 python tools/pnp_timing.py > gpurun_out/pnp_timing.json 2> gpurun_out/pnp_timing.err; echo "timing rc=$?"
cat gpurun_out/pnp_timing.json; tail -5 gpurun_out/pnp_timing.err
