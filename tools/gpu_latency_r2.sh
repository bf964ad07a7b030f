#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/latency.py 2>&1 | tail -1
MSLAM_B200_LATENCY_PATH=0 timeout 200 python tools/latency.py 2>&1 | tail -1
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
