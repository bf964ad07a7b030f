#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --workload c5 --steps 5 --warmup 3 > gpurun_out/bench_c5.log 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"; tail -3 gpurun_out/bench_c5.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_c5.log").read().strip().splitlines()[-1])
print(d["value"], d["roofline"]["frac_of_tmem_read_bound"], d["roofline"]["int_pipe_path"])
PY
export MSLAM_BENCH_SKIP_INT_PIPE=1 MSLAM_BENCH_KF_PER_GPU=1250
BENCH_ARGS="--workload c5" NCU_KERNELS="k_knn2_mma|k_expand_pm1" NCU_SKIP=4 NCU_COUNT=4 bash tools/gpu_profile.sh > gpurun_out/profile_c5.log 2>&1
tail -4 gpurun_out/profile_c5.log
mv gpurun_out/prof.ncu-rep gpurun_out/prof_c5.ncu-rep; mv gpurun_out/launches.csv gpurun_out/launches_c5.csv
