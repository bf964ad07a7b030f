#!/bin/bash
mkdir -p gpurun_out
MSLAM_KNN_MMA=2 timeout 600 python -m pytest tests -m gpu -q -x --timeout 200 -k "match or tie or knn or frontend or db_ or keyframe" > gpurun_out/pytest_mma.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_mma.log
tail -15 gpurun_out/pytest_mma.log
grep -q "pytest rc=0" gpurun_out/pytest_mma.log && bash tools/gpu_mma2.sh
