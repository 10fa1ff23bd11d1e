#!/bin/bash
# Round 2, one GPU: default bench line with one and with two jobs in flight; cfg3 with two; GPU suite.
tag=${1:-r02y}
mkdir -p gpurun_out
LIME_BENCH_INFLIGHT=1 timeout 600 python bench.py --no-cpu --no-e2e > gpurun_out/${tag}_bench_inflight1.json 2> gpurun_out/${tag}_bench_inflight1.err; echo "bench inflight1 rc=$?"; tail -3 gpurun_out/${tag}_bench_inflight1.err | cut -c1-300
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err | cut -c1-300
timeout 900 python bench.py --config cfg3 --no-e2e --steps 5 > gpurun_out/${tag}_bench_cfg3_n1.json 2> gpurun_out/${tag}_bench_cfg3_n1.err; echo "bench cfg3 rc=$?"; tail -3 gpurun_out/${tag}_bench_cfg3_n1.err | cut -c1-300
timeout 900 python bench.py --config cfg4 --no-e2e --steps 5 > gpurun_out/${tag}_bench_cfg4_n1.json 2> gpurun_out/${tag}_bench_cfg4_n1.err; echo "bench cfg4 rc=$?"; tail -3 gpurun_out/${tag}_bench_cfg4_n1.err | cut -c1-300
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
python - <<PY
import json, glob
for f in sorted(glob.glob('gpurun_out/${tag}_bench*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        r = d['roofline']
        print(f, 'value %.4g ms/step %.3f (instrumented %.3f) frac %.3f' % (d['value'], d['ms_per_step'], r['ms_per_step_with_kernel_events'], r['frac']), d['config']['parallelism'], 'parity', d['parity'].get('checked'), 'e2e', (d.get('e2e') or {}).get('value'))
    except Exception as e:
        print(f, 'unreadable:', e)
PY
