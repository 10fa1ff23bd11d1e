#!/bin/bash
# Round 2, final one-GPU call: smoke(), GPU suite, default bench line, reference arm, launch list of the bench command.
tag=${1:-r02ab}
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 gpurun_out/${tag}_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err | cut -c1-300
timeout 600 python bench.py --impl reference > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "reference arm rc=$?"
K="regex:stream_kernel|bytesum_kernel|warp_kernel|block_kernel|count_kernel|emit_kernel|choose_kernel|choose_sparse_kernel|choose_wide_kernel|classify_kernel|narrow_kernel|route_kernel|apply_kernel|chunk_scan|expand"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
python - <<PY
import json
for f in ('gpurun_out/${tag}_bench.json', 'gpurun_out/${tag}_ref.json'):
    d = json.loads([l for l in open(f) if l.startswith('{')][-1])
    r = d.get('roofline') or {}
    print(f, 'value %.4g ms/step %.3f frac %s' % (d['value'], d['ms_per_step'], r.get('frac')), 'cpu', (d.get('cpu_baseline') or {}).get('value'), 'parity', (d.get('parity') or {}).get('checked'), 'e2e', (d.get('e2e') or {}).get('modes'))
PY
