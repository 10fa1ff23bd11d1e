#!/bin/bash
# Round 2: route_kernel tile size (threads x 8 records) and apply row-block size on N GPUs, cfg2, no parity / e2e legs.
#   gpurun --gpus N --timeout 900 -- 'bash tools/gpu_r2_q.sh <tag> <N> "<threads>:<sub_mb>" ...'
tag=${1:-r02q}; N=${2:-8}; shift 2
mkdir -p gpurun_out
port=29740
for v in "$@"; do
  th=${v%%:*}; mb=${v##*:}; port=$((port+1))
  LIME_XCHG_ROUTE_THREADS=$th LIME_XCHG_SUB_MB=$mb timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $N --config cfg2 --no-e2e --no-parity --steps 8 > gpurun_out/${tag}_n${N}_t${th}_s${mb}.json 2> gpurun_out/${tag}_n${N}_t${th}_s${mb}.err
  echo "threads=$th sub_mb=$mb rc=$?"; tail -2 gpurun_out/${tag}_n${N}_t${th}_s${mb}.err | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads([l for l in open('gpurun_out/${tag}_n${N}_t${th}_s${mb}.json') if l.startswith('{')][-1])
    r = d['roofline']; x = d['exchange']
    print('threads=$th sub_mb=$mb n=$N: value %.4g ms/step %.3f k12 %s' % (d['value'], d['ms_per_step'], r['k12_ms_per_launch_by_rank'][:2]))
    print('   kernels', {k: round(v, 3) for k, v in r['kernels_ms_per_step_rank0'].items()})
except Exception as e:
    print('bench unreadable:', e)
PY
done
