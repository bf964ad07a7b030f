#!/bin/bash
# end-of-round N=8 bench lines (one node, torchrun, NCCL)
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $R --master-port 29601 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/n8_c2.json 2> gpurun_out/n8_c2.err; echo c2 rc=$?
timeout 200 $R --master-port 29602 bench.py --gpus 8 --workload pnp --no-cpu-baseline > gpurun_out/n8_pnp.json 2> gpurun_out/n8_pnp.err; echo pnp rc=$?
timeout 300 $R --master-port 29603 bench.py --gpus 8 --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/n8_c5.json 2> gpurun_out/n8_c5.err; echo c5 rc=$?
python - <<'P'
import json
for f in ["c2","pnp","c5"]:
    try:
        d=json.loads(open(f"gpurun_out/n8_{f}.json").read().strip().splitlines()[-1]); print(f, d["n_gpus"], d["value"], d["unit"], d["e2e"]["value"], d["ms_per_step"])
    except Exception as e: print(f, "ERR", e)
P
