#!/bin/bash
# GPU parity tests + keyframe-DB bench (c5) at N GPUs (default 1)
mkdir -p gpurun_out
N=${N:-1}
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
if [ "$N" = "1" ]; then
  timeout 900 python bench.py --workload c5 --steps ${STEPS:-5} --warmup 3 > gpurun_out/bench_c5.log 2> gpurun_out/bench_c5.err
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload c5 --steps ${STEPS:-5} --warmup 3 > gpurun_out/bench_c5_n$N.log 2> gpurun_out/bench_c5_n$N.err
fi
echo "bench rc=$?"
tail -2 gpurun_out/bench_c5*.log; tail -5 gpurun_out/bench_c5*.err
