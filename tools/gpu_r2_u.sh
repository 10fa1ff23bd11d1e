#!/bin/bash
# Round 2, one GPU: the no-allocation entry point alone (fresh process) and after the whole GPU suite, then the
# default bench line, the reference arm and the config-5 sweep.
tag=${1:-r02u}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${tag}_pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -15 gpurun_out/${tag}_pytest_no_alloc_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -5 gpurun_out/${tag}_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err | cut -c1-300
timeout 600 python bench.py --impl reference > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "reference arm rc=$?"
for a in 12 14 16 20 24; do
  timeout 900 python bench.py --n 1000000000 --alpha $a --no-e2e --steps 5 > gpurun_out/${tag}_sweep_a$a.json 2> gpurun_out/${tag}_sweep_a$a.err
  echo "sweep alpha=$a rc=$?"; tail -2 gpurun_out/${tag}_sweep_a$a.err | cut -c1-300
done
python - <<PY
import json, glob
for f in sorted(glob.glob('gpurun_out/${tag}_sweep_a*.json')) + ['gpurun_out/${tag}_bench.json', 'gpurun_out/${tag}_ref.json']:
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        r = d.get('roofline') or {}
        print(f, 'value %.4g ms/step %.3f frac %s' % (d['value'], d['ms_per_step'], r.get('frac')), {k: round(v, 3) for k, v in r.get('kernels_ms_per_step', {}).items()},
              'cpu', (d.get('cpu_baseline') or {}).get('value'), 'parity', (d.get('parity') or {}).get('checked'), 'e2e', (d.get('e2e') or {}).get('modes'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
