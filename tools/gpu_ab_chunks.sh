for V in MSLAM_B200_CHUNKS=1 MSLAM_B200_CHUNKS=2 MSLAM_B200_CHUNKS=4; do
  env $V python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab.log 2> gpurun_out/ab.err
  python -c "
import json;d=json.loads(open('gpurun_out/ab.log').readline());e=d['e2e'];print('$V', 'pipelined %.0f (%.3f ms)  sync %.0f (%.3f ms)'%(e['value'],e['ms_per_step'],e['synchronous']['value'],e['synchronous']['ms_per_step']))"
done
