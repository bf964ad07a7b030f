#!/bin/bash
# copy the outputs of tools/gpu_final_r2.sh from gpurun_out/ into profiles/ under a tag:  tools/collect_final.sh r2_v7
tag=$1
for w in c2 ref c3 c4 c5 dist pnp; do python - $w $tag <<'P'
import sys
w,tag=sys.argv[1:3]
txt=open(f'gpurun_out/f_{w}.json').read().strip().splitlines()[-1]
open(f'profiles/{tag}_bench_{w}.json','w').write(txt[txt.index('{'):]+"\n")
P
done
cp gpurun_out/f_pytest.log profiles/${tag}_pytest_gpu.log; cp gpurun_out/f_launches.csv profiles/${tag}_launches.csv
python tools/ncu_traffic.py gpurun_out/f_traffic.csv $(git rev-parse --short HEAD) > profiles/ncu_traffic.json
python tools/ncu_summary.py launches gpurun_out/f_launches.csv > profiles/${tag}_launch_summary.txt 2>&1
for w in c2 ref c3 c4 c5 dist pnp; do python - $w $tag <<'P'
import json,sys
w,tag=sys.argv[1:3]
d=json.load(open(f'profiles/{tag}_bench_{w}.json')); e=d.get('e2e',{})
print(w, round(d['value'],1), d['unit'], 'ms/step', round(d['ms_per_step'],4), 'e2e', round(e.get('value',0),1), 'frac', round(d.get('roofline',{}).get('frac',0) or 0,4), 'cpu', (d.get('cpu_baseline') or {}).get('value'))
if w=='c2': print('  plugin', e.get('plugin_single_frame',{}).get('value'), e.get('plugin_batched',{}).get('value'), 'db', d.get('db',{}).get('value'), 'c3', d.get('c3',{}).get('value'), {k:round(v['ms_per_step'],3) for k,v in d['roofline']['stages'].items()})
P
done
head -12 profiles/${tag}_launch_summary.txt
