#!/bin/bash
# Round 2, multi-GPU call for the side-stream exchange: loopback tests, torchrun parity, bench lines with 5 / 6 / 7 / all
# CTA slots of an SM given to the scoring kernel.
#   gpurun --gpus N --timeout 1800 -- 'bash tools/gpu_r2_p.sh <tag> <N> [ctas...]'
tag=${1:-r02p}; N=${2:-2}; shift 2
ctas=${@:-6 7 0}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "xchg or deferred" > gpurun_out/${tag}_pytest_xchg.log 2>&1; echo "pytest xchg rc=$?"; tail -4 gpurun_out/${tag}_pytest_xchg.log
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/${tag}_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -8 gpurun_out/${tag}_pytest_multi.log
port=29640
for c in $ctas; do
  port=$((port+1))
  LIME_XCHG_K12_CTAS=$c timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $N --config cfg2 --no-e2e > gpurun_out/${tag}_bench_cfg2_n${N}_c$c.json 2> gpurun_out/${tag}_bench_cfg2_n${N}_c$c.err
  echo "bench ctas=$c rc=$?"; tail -3 gpurun_out/${tag}_bench_cfg2_n${N}_c$c.err | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads([l for l in open('gpurun_out/${tag}_bench_cfg2_n${N}_c$c.json') if l.startswith('{')][-1])
    r = d['roofline']; x = d['exchange']
    print('ctas=$c n=$N: value %.4g ms/step %.3f (profiled run %.3f) k12 %s parity %s route %s' % (d['value'], d['ms_per_step'], x.get('ms_per_step_of_the_profiled_run_rank0', 0),
          r['k12_ms_per_launch_by_rank'], d['parity'].get('checked'), x['route']))
    print('   kernels', {k: round(v, 3) for k, v in r['kernels_ms_per_step_rank0'].items()})
except Exception as e:
    print('bench unreadable:', e)
PY
done
