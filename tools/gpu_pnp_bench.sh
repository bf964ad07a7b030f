#!/bin/bash
# bench line + ncu launch list + one full capture of k_pnp_ransac
mkdir -p gpurun_out
timeout 300 python bench.py --workload pnp --steps 20 --warmup 5 > gpurun_out/bench_pnp.log 2> gpurun_out/bench_pnp.err; echo "bench rc=$?"
tail -1 gpurun_out/bench_pnp.log; tail -3 gpurun_out/bench_pnp.err
CMD="python bench.py --workload pnp --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/pnp_launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_pnp_ransac -s 3 -c 2 -o gpurun_out/pnp_prof $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
timeout 120 python bench.py --impl reference --workload pnp --steps 2 --warmup 1 > gpurun_out/bench_pnp_ref.log 2>&1; echo "ref rc=$?"
tail -1 gpurun_out/bench_pnp_ref.log
