#!/bin/bash
# Round 2, one GPU: BASELINE config 5 mini-sweep (10^9 symbols per strand, alpha 12..24), ncu of the block-path
# kernel at alpha 12, and the final launch list + K12 capture of the default bench command.
#   gpurun --timeout 2400 -- 'bash tools/gpu_r2_s.sh <tag>'
tag=${1:-r02s}
mkdir -p gpurun_out
for a in 12 16 20 24; do
  timeout 900 python bench.py --n 1000000000 --alpha $a --no-e2e --steps 5 > gpurun_out/${tag}_sweep_a$a.json 2> gpurun_out/${tag}_sweep_a$a.err
  echo "sweep alpha=$a rc=$?"; tail -2 gpurun_out/${tag}_sweep_a$a.err | cut -c1-300
done
python - <<PY
import json, glob
for f in sorted(glob.glob('gpurun_out/${tag}_sweep_a*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
        r = d['roofline']; st = d['config'].get('strand_stats', [{}])[0]
        print(f, 'value %.4g ms/step %.3f frac %.3f' % (d['value'], d['ms_per_step'], r['frac']), {k: round(v, 3) for k, v in r.get('kernels_ms_per_step', {}).items()},
              'cpu', (d.get('cpu_baseline') or {}).get('value'), 'parity', d.get('parity', {}).get('checked'), st)
    except Exception as e:
        print(f, 'unreadable:', e)
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:block_kernel -c 1 -o gpurun_out/${tag}_k2block_a12 \
    python bench.py --n 1000000000 --alpha 12 --no-e2e --no-cpu --steps 1 --warmup 3 > gpurun_out/${tag}_ncu_k2block.log 2>&1; echo "block capture rc=$?"
# the default bench command: plain run, launch list, K12 capture
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "reference arm rc=$?"; cat gpurun_out/${tag}_ref.json | cut -c1-400
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_list.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 2 -c 1 -o gpurun_out/${tag}_k12 \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_k12.log 2>&1; echo "k12 capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:choose_wide -c 1 -o gpurun_out/${tag}_k3wide \
    python bench.py --config cfg3 --n 400000000 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/${tag}_ncu_k3wide.log 2>&1; echo "wide choose capture rc=$?"
python - <<PY
import json
d = json.loads([l for l in open('gpurun_out/${tag}_bench.json') if l.startswith('{')][-1])
r = d['roofline']
print('bench: value %.4g ms/step %.3f frac %.3f e2e %s cpu %s' % (d['value'], d['ms_per_step'], r['frac'], (d.get('e2e') or {}).get('value'), (d.get('cpu_baseline') or {}).get('value')))
PY
