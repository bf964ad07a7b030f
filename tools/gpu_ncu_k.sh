#!/bin/bash
# ncu --set full of selected kernels on a short bench run:  NCU_KERNELS=regex NCU_SKIP=n NCU_COUNT=n BENCH_ARGS=...
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline ${BENCH_ARGS}"
$CMD > gpurun_out/plain.log 2> gpurun_out/plain.err || exit 1
ncu --set full --clock-control none --import-source on -k "regex:${NCU_KERNELS}" -s ${NCU_SKIP:-0} -c ${NCU_COUNT:-10} -f -o gpurun_out/${NCU_OUT:-prof_k} $CMD > gpurun_out/ncu_k.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_k.log
