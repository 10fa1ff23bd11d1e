#!/bin/bash
# bench (+ optional ncu capture) of several builds of the library: gpu_variants.sh <tag> <ncu:0|1> <suffix...>   ("-" = the default build)
tag=$1; ncu=$2; shift 2
mkdir -p gpurun_out
for v in "$@"; do
  [ "$v" = "-" ] && v=""
  export LIME_B200_LIB=$PWD/lime_b200/liblime_b200$v.so
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/${tag}_bench$v.json 2> gpurun_out/${tag}_bench$v.err; echo "bench$v rc=$?"
  python - <<PY
import json
try:
    d = json.load(open('gpurun_out/${tag}_bench$v.json')); r = d['roofline']
    print('lib$v: ms/step %.3f frac %.3f %s' % (d['ms_per_step'], r['frac'], {k: round(x, 3) for k, x in r['kernels_ms_per_step'].items()}))
except Exception as e:
    print('bench$v unreadable:', e)
PY
  if [ "$ncu" = 1 ]; then
    timeout 300 ncu --set full --clock-control none --import-source on -k "regex:stream_kernel" -s 2 -c 1 -o gpurun_out/${tag}_k12$v python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_k12$v.log 2>&1
    echo "k12$v capture rc=$?"
  fi
done
