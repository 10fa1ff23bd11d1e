#!/bin/bash
# compile-time experiments of the stream kernel (LIME_K12_EXP bits, see k12_stream.cu): rebuilds the
# library on the GPU box per variant, runs the parity tests of the fused path and a short bench, and
# puts the default build back.  usage: gpu_exp.sh 1 2 4 8 12 15   (1: row address per read leader, 2: arithmetic A,C,G,T class, 4: packed stream element, 8: phase C visits non-empty words only; sums combine)
mkdir -p gpurun_out
for v in ${@:-1 2 4 8 12 15} 0; do
  touch lime_b200/csrc/k12_stream.cu
  make -s -C lime_b200/csrc EXTRA=-DLIME_K12_EXP=$v || { echo "build $v failed"; continue; }
  if [ "$v" != 0 ]; then
    timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "steps_1_2 or chunk_boundaries or adversarial or medium" > gpurun_out/exp_${v}_pytest.log 2>&1; echo "exp=$v pytest rc=$?"
  fi
  python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/exp_$v.json 2> gpurun_out/exp_$v.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/exp_$v.json')); r=d['roofline']
    print('exp=$v', 'ms/step=%.3f'%d['ms_per_step'], {k:round(x,3) for k,x in r['kernels_ms_per_step'].items()})
except Exception as e:
    print('exp=$v failed', e)
PY
done
