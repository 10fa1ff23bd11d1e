#!/bin/bash
tag=${1:-r02ag}
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/${tag}_bench.err | cut -c1-400
python - <<PY
import json
d = json.loads([l for l in open('gpurun_out/${tag}_bench.json') if l.startswith('{')][-1])
print('value %.4g ms/step %.3f frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']), 'cpu', (d.get('cpu_baseline') or {}).get('value'), 'parity', d['parity'].get('checked'), 'e2e', (d.get('e2e') or {}).get('modes'))
PY
