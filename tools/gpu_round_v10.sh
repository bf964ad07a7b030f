#!/bin/bash
# full measurement pass: GPU tests, smoke, the three bench lines (c2, dist, c5) with CPU baselines
bash tools/gpu_round.sh
timeout 600 python bench.py --workload dist --steps 20 --warmup 5 > gpurun_out/bench_dist.log 2> gpurun_out/bench_dist.err; echo "dist rc=$?"
timeout 900 python bench.py --workload c5 --steps 5 --warmup 3 > gpurun_out/bench_c5.log 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"
python - <<'PY'
import json
for n in ("bench", "bench_dist", "bench_c5"):
    try:
        d = json.loads(open(f"gpurun_out/{n}.log").read().strip().splitlines()[-1])
        print(n, d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel"], round(d["roofline"]["frac"], 4), d["clocks"])
    except Exception as e:
        print(n, "failed", e)
PY
