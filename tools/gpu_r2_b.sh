#!/bin/bash
# Round 2, call B: GPU suite, then the stream kernel at 8 and at 7 CTAs per SM (LIME_B200_LIB picks the build),
# each with a short bench and a full ncu capture.
tag=${1:-r02b}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -12 gpurun_out/${tag}_pytest_gpu.log
CMD="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu"
for v in "" _s; do
  export LIME_B200_LIB=$PWD/lime_b200/liblime_b200$v.so
  timeout 300 $CMD > gpurun_out/${tag}_bench$v.json 2> gpurun_out/${tag}_bench$v.err; echo "bench$v rc=$?"
  python - <<PY
import json
try:
    d = json.load(open('gpurun_out/${tag}_bench$v.json')); r = d['roofline']
    print('lib$v: ms/step %.3f frac %.3f %s' % (d['ms_per_step'], r['frac'], {k: round(x, 3) for k, x in r['kernels_ms_per_step'].items()}))
except Exception as e:
    print('bench$v unreadable:', e)
PY
  timeout 300 ncu --set full --clock-control none --import-source on -k "regex:stream_kernel" -s 2 -c 1 -o gpurun_out/${tag}_k12$v python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_k12$v.log 2>&1
  echo "k12$v capture rc=$?"
done
unset LIME_B200_LIB
timeout 300 python tools/k12_modes.py > gpurun_out/${tag}_modes.log 2>&1; echo "modes rc=$?"; cat gpurun_out/${tag}_modes.log | cut -c1-400
