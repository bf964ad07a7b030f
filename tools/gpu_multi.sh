#!/bin/bash
# multi-GPU bench lines (c2 replicas, c5 sharded DB) at N GPUs:  N=2 bash tools/gpu_multi.sh
mkdir -p gpurun_out
N=${N:-2}
nvidia-smi -L > gpurun_out/smi_multi.txt
for WL in ${WLS:-c2 c5}; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $WL --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline > gpurun_out/bench_${WL}_n$N.log 2> gpurun_out/bench_${WL}_n$N.err
  echo "$WL N=$N rc=$?"; tail -1 gpurun_out/bench_${WL}_n$N.log | cut -c1-400; tail -3 gpurun_out/bench_${WL}_n$N.err
done
