#!/bin/bash
# round 2: ncu launch list of the c2 step + one --set full capture of its kernels (B200_PROFILING.md recipe), then the same
# for the keyframe-database scoring kernel.  Every ncu pass runs only after the same command has exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sub"
timeout 200 $CMD > gpurun_out/p_plain.log 2> gpurun_out/p_plain.err || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/p_list.log 2>&1
echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_fast|k_orient_describe|k_blur|k_knn2|k_select|k_resize|k_gray|k_expand|k_ratio" -s 70 -c 30 -f -o gpurun_out/r2_prof $CMD > gpurun_out/p_full.log 2>&1
echo "full capture rc=$?"; tail -2 gpurun_out/p_full.log
DB="python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline"
MSLAM_BENCH_SKIP_INT_PIPE=1 MSLAM_BENCH_KF_PER_GPU=2500 timeout 200 $DB > gpurun_out/p_db.log 2> gpurun_out/p_db.err || exit 1
MSLAM_BENCH_SKIP_INT_PIPE=1 MSLAM_BENCH_KF_PER_GPU=2500 timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_knn2_mma" -s 3 -c 2 -f -o gpurun_out/r2_prof_db $DB > gpurun_out/p_dbfull.log 2>&1
echo "db capture rc=$?"; tail -2 gpurun_out/p_dbfull.log
ls -la gpurun_out | head -30
