#!/bin/bash
# Round 2: one vs two jobs in flight on N GPUs (LIME_BENCH_INFLIGHT), cfg2 (+ configs given), parity legs on.
tag=${1:-r02x}; N=${2:-2}; shift 2
cfgs=${@:-cfg2}
mkdir -p gpurun_out
port=29940
for cfg in $cfgs; do
 for inf in 1 2; do
  port=$((port+1))
  extra="--no-e2e"; [ "$cfg" != cfg2 ] && extra="--no-e2e --steps 6"
  LIME_BENCH_INFLIGHT=$inf timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $N --config $cfg $extra > gpurun_out/${tag}_${cfg}_n${N}_inflight$inf.json 2> gpurun_out/${tag}_${cfg}_n${N}_inflight$inf.err
  echo "$cfg inflight=$inf rc=$?"; tail -2 gpurun_out/${tag}_${cfg}_n${N}_inflight$inf.err | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads([l for l in open('gpurun_out/${tag}_${cfg}_n${N}_inflight$inf.json') if l.startswith('{')][-1])
    r = d['roofline']; x = d['exchange']
    print('$cfg inflight=$inf n=$N: value %.4g ms/step %.3f (profiled %.3f) k12 %s parity %s route %s' % (d['value'], d['ms_per_step'], x.get('ms_per_step_of_the_profiled_run_rank0', 0), r['k12_ms_per_launch_by_rank'][:2], d['parity'].get('checked'), x['route']))
    print('   kernels', {k: round(v, 3) for k, v in r['kernels_ms_per_step_rank0'].items()}, 'launches', d['gpu_launches'])
except Exception as e:
    print('bench unreadable:', e)
PY
 done
done
