#!/bin/bash
# Round 2, one-GPU call: GPU suite, default bench line, the north-star shapes on one GPU.
#   gpurun --timeout 1800 -- 'bash tools/gpu_r2_o.sh <tag> [configs...]'
tag=${1:-r02o}; shift
cfgs=${@:-cfg3 cfg4}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -5 gpurun_out/${tag}_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err | cut -c1-300
for cfg in $cfgs; do
  timeout 900 python bench.py --config $cfg --no-e2e --steps 5 > gpurun_out/${tag}_bench_${cfg}_n1.json 2> gpurun_out/${tag}_bench_${cfg}_n1.err
  echo "bench $cfg rc=$?"; tail -3 gpurun_out/${tag}_bench_${cfg}_n1.err | cut -c1-300
done
python - <<PY
import json, glob
for f in sorted(glob.glob('gpurun_out/${tag}_bench*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        r = d['roofline']
        print(f, 'value %.4g ms/step %.3f frac %.3f' % (d['value'], d['ms_per_step'], r['frac']), {k: round(v, 3) for k, v in r.get('kernels_ms_per_step', {}).items()}, 'e2e', (d.get('e2e') or {}).get('value'), 'parity', d.get('parity', {}).get('checked'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
