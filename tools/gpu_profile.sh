#!/bin/bash
# ncu launch list + one --set full capture of the hot kernels, on a short bench run (B200_PROFILING.md recipe)
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline ${BENCH_ARGS}"
$CMD > gpurun_out/plain.log 2> gpurun_out/plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2.log 2> gpurun_out/plain2.err &&
ncu --set full --clock-control none --import-source on -k "regex:${NCU_KERNELS:-k_fast|k_orient_describe|k_blur|k_knn2|k_select|k_resize|k_gray}" -s ${NCU_SKIP:-81} -c ${NCU_COUNT:-27} -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out
