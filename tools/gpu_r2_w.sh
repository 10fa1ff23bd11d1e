#!/bin/bash
# Round 2, one GPU: the narrowed-transfer tests and the default bench line (e2e: lcp bits + da as 3 bytes; LIME_PACK_DA=0 for comparison)
tag=${1:-r02w}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zz_gpu_narrow_load.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${tag}_pytest_narrow.log 2>&1; echo "pytest narrow rc=$?"; tail -6 gpurun_out/${tag}_pytest_narrow.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err | cut -c1-300
LIME_PACK_DA=0 timeout 600 python bench.py --no-cpu > gpurun_out/${tag}_bench_da4.json 2> gpurun_out/${tag}_bench_da4.err; echo "bench da4 rc=$?"
python - <<PY
import json
for f in ('gpurun_out/${tag}_bench.json', 'gpurun_out/${tag}_bench_da4.json'):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value %.4g ms/step %.3f frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']), 'cpu', (d.get('cpu_baseline') or {}).get('value'), 'e2e', (d.get('e2e') or {}).get('modes'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
