#!/bin/bash
# A/B runs of bench.py under different env settings:  VARS="MSLAM_FAST_MINB=3 MSLAM_FAST_MINB=4" bash tools/gpu_ab.sh
mkdir -p gpurun_out
for V in $VARS; do
  env $V python bench.py --steps 20 --warmup 5 --no-cpu-baseline ${BENCH_ARGS} > gpurun_out/ab_$V.log 2> gpurun_out/ab_$V.err
  python - "$V" <<'P'
import json,sys
v=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/ab_{v}.log').read().strip().splitlines()[-1])
    st=d['roofline']['stages']
    print(v, 'fps %.0f e2e %.0f |'%(d['value'], d['e2e']['value']), ' '.join('%s %.4f'%(k,x['ms_per_step']) for k,x in st.items()))
except Exception as e:
    print(v, 'FAILED', e)
P
done
