#!/bin/bash
# one --set full capture (with source) of the big k_fast launch and of k_blur / k_orient_describe of a c2 step
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sub"
timeout 200 $CMD > gpurun_out/p_plain.log 2> gpurun_out/p_plain.err || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_fast|k_blur|k_orient_describe" -s 30 -c 8 -f -o gpurun_out/r2_prof_fast $CMD > gpurun_out/p_full.log 2>&1
echo "full capture rc=$?"; tail -2 gpurun_out/p_full.log
