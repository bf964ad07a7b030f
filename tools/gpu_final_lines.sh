python bench.py --workload dist --no-cpu-baseline 2>/dev/null > gpurun_out/final_dist.json
python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/final_c5.json
python bench.py --workload c3 --no-cpu-baseline 2>/dev/null > gpurun_out/final_c3.json
python bench.py --workload c4 --steps 10 --no-cpu-baseline 2>/dev/null > gpurun_out/final_c4.json
python bench.py 2>/dev/null > gpurun_out/final_c2.json
python - <<'P'
import json
for f in ["dist","c5","c3","c4","c2"]:
    try:
        d=json.loads(open(f"gpurun_out/final_{f}.json").read().strip().splitlines()[-1]); print(f, d["value"], d["unit"], d["e2e"]["value"], d["ms_per_step"])
    except Exception as e: print(f, "ERR", e)
P
