#!/bin/bash
# Round 2: the exchange on N GPUs with and without the side stream (LIME_XCHG_TAIL), cfg2; loopback tests first.
tag=${1:-r02r}; N=${2:-2}; shift 2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "xchg or deferred" > gpurun_out/${tag}_pytest_xchg.log 2>&1; echo "pytest xchg rc=$?"; tail -4 gpurun_out/${tag}_pytest_xchg.log
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/${tag}_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -8 gpurun_out/${tag}_pytest_multi.log
port=29840
for tail in 0 1; do
  port=$((port+1))
  LIME_XCHG_TAIL=$tail timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $N --config cfg2 --no-e2e --no-parity --steps 8 > gpurun_out/${tag}_n${N}_tail$tail.json 2> gpurun_out/${tag}_n${N}_tail$tail.err
  echo "tail=$tail rc=$?"; tail -2 gpurun_out/${tag}_n${N}_tail$tail.err | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads([l for l in open('gpurun_out/${tag}_n${N}_tail$tail.json') if l.startswith('{')][-1])
    r = d['roofline']; x = d['exchange']
    print('tail=$tail n=$N: value %.4g ms/step %.3f k12 %s' % (d['value'], d['ms_per_step'], r['k12_ms_per_launch_by_rank'][:2]))
    print('   kernels', {k: round(v, 3) for k, v in r['kernels_ms_per_step_rank0'].items()})
except Exception as e:
    print('bench unreadable:', e)
PY
done
