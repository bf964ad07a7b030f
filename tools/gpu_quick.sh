#!/bin/bash
# quick A/B: detector parity tests + the c2 line (device value, stages); QUICK_VARIANTS="ENV=.. ENV2=.." adds runs under those settings
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_oracle_vs_cv2.py tests/test_reference_build.py -m gpu -x -q 2>&1 | tail -3
run() {
timeout 300 env $1 python bench.py --steps 20 --warmup 3 --no-sub --no-cpu-baseline > gpurun_out/q_c2.json 2> gpurun_out/q_c2.err; tail -1 gpurun_out/q_c2.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/q_c2.json').read().strip().splitlines()[-1])
print(round(d['value']), d['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in d['roofline']['stages'].items()}, d['natural_frames']['frames_per_s'])
P
}
run X=1; run X=1
for v in $QUICK_VARIANTS; do echo "== $v"; run $v; done
