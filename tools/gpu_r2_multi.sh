#!/bin/bash
# Round 2, multi-GPU call: torchrun parity test on all GPUs of the box, then bench lines.
#   gpurun --gpus N --timeout 1800 -- 'bash tools/gpu_r2_multi.sh <tag> <N> [configs...]'
tag=${1:-r02m}; N=${2:-2}; shift 2
cfgs=${@:-cfg2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
nvidia-smi topo -m 2>/dev/null | head -12 > gpurun_out/${tag}_topo.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/${tag}_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -8 gpurun_out/${tag}_pytest_multi.log
port=29540
for cfg in $cfgs; do
  port=$((port+1))
  extra=""; [ "$cfg" != cfg2 ] && extra="--no-e2e --steps 5"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $N --config $cfg $extra > gpurun_out/${tag}_bench_${cfg}_n$N.json 2> gpurun_out/${tag}_bench_${cfg}_n$N.err
  echo "bench $cfg n=$N rc=$?"; tail -4 gpurun_out/${tag}_bench_${cfg}_n$N.err | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads([l for l in open('gpurun_out/${tag}_bench_${cfg}_n$N.json') if l.startswith('{')][-1])
    r = d['roofline']; x = d['exchange']
    print('$cfg n=$N: value %.4g ms/step %.3f job_frac %.3f route %s k12 %s phases %s e2e %s parity %s' % (d['value'], d['ms_per_step'], r['job_frac_of_aggregate_peak'],
          x['route'], r['k12_ms_per_launch_by_rank'], {k: round(v, 2) for k, v in x['phases_ms_max_over_ranks'].items()}, (d.get('e2e') or {}).get('value'), d['parity'].get('checked')))
    print('   kernels', {k: round(v, 3) for k, v in r['kernels_ms_per_step_rank0'].items()}, 'nvlink GB/s', x.get('route_kernel_nvlink_GBps_rank0'))
except Exception as e:
    print('bench $cfg unreadable:', e)
PY
done
