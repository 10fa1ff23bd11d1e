#!/bin/bash
# Round 2, call A: the whole GPU suite on the rewritten stream kernel (TMA staging, packed stream, 8 CTAs/SM,
# LCP8 / flat variants), the record exchange in loopback, the deferred strands; then the bench line, the ncu
# launch list and one full capture of the stream kernel.
#   gpurun --timeout 1500 -- 'bash tools/gpu_r2_a.sh <tag>'
tag=${1:-r02a}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -15 gpurun_out/${tag}_pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench_n1.err
python - <<PY
import json
try:
    d = json.load(open('gpurun_out/${tag}_bench_n1.json'))
    r = d['roofline']
    print('value %.4g ms/step %.3f  e2e %.4g  frac %.3f  %s' % (d['value'], d['ms_per_step'], (d.get('e2e') or {}).get('value', 0), r['frac'],
          {k: round(x, 3) for k, x in r['kernels_ms_per_step'].items()}))
except Exception as e:
    print('bench line unreadable:', e)
PY
K="regex:stream_kernel|bytesum_kernel|warp_kernel|block_kernel|count_kernel|emit_kernel|choose_kernel|choose_sparse_kernel|classify_kernel|narrow_kernel|route_kernel|apply_kernel|chunk_scan"
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:stream_kernel" -s 2 -c 1 -o gpurun_out/${tag}_k12 $CMD > gpurun_out/${tag}_ncu_k12.log 2>&1
echo "k12 capture rc=$?"
