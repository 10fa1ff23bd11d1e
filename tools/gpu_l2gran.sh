#!/bin/bash
for g in 128 64 32; do
  LIME_L2_GRAN=$g python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/l2_$g.json 2> gpurun_out/l2_$g.err
  grep "L2 fetch" gpurun_out/l2_$g.err | head -1
  python -c "
import json; d=json.load(open('gpurun_out/l2_$g.json')); print('gran $g', d['ms_per_step'], d['roofline']['kernels_ms_per_step'])"
done
