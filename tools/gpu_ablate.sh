#!/bin/bash
# ablation of the fused kernel through the LIME_DEBUG switches (diagnostic; results are wrong with dbg != 0)
# bit0: no global scalars, bit1: no matrix atomics, bit2: stop before phase D, bit3: stop before phase C,
# bit4: atomics without return value
mkdir -p gpurun_out
for d in ${@:-0 2 4 8}; do
  LIME_DEBUG=$d python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu > gpurun_out/abl_$d.json 2> gpurun_out/abl_$d.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/abl_$d.json')); r=d['roofline']
    print('dbg=$d', 'ms/step=%.3f'%d['ms_per_step'], {k:round(v,3) for k,v in r['kernels_ms_per_step'].items()}, 'frac=%.3f'%r['frac'])
except Exception as e:
    print('dbg=$d failed', e)
PY
done
