#!/bin/bash
# end-of-round evidence: GPU tests, bench lines of every workload WITH their CPU baselines, reference arm, launch list, DRAM traffic
mkdir -p gpurun_out
timeout 200 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -1
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/f_pytest.log; tail -2 gpurun_out/f_pytest.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/f_c2.json 2> gpurun_out/f_c2.err; tail -1 gpurun_out/f_c2.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/f_ref.json 2> gpurun_out/f_ref.err
for w in c3 c4 c5 dist pnp; do timeout 300 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/f_$w.json 2> gpurun_out/f_$w.err; echo "$w rc=$?"; done
CMD="python bench.py --serial --steps 1 --warmup 3 --no-sub --no-cpu-baseline"
timeout 200 $CMD > /dev/null 2>&1 && timeout 400 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 90 --csv --log-file gpurun_out/f_traffic.csv $CMD > gpurun_out/f_traffic.log 2>&1
echo "traffic rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-sub --no-cpu-baseline"
timeout 200 $CMD > /dev/null 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/f_launches.csv $CMD > gpurun_out/f_list.log 2>&1
echo "launch list rc=$?"
