#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
python - <<'P' > gpurun_out/int_peaks.log 2>&1
import __graft_entry__ as g
pkg = g.load_package()
with pkg.Context() as c:
    for _ in range(3):
        print(c.measure_int_peaks())
P
cat gpurun_out/int_peaks.log
