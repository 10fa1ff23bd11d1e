#!/bin/bash
# Round 2: the memory guard of the second context: cfg4 on 2 GPUs (does not fit twice) and cfg2 (fits)
tag=${1:-r02ai}; N=2
mkdir -p gpurun_out
for cfg in cfg4 cfg2; do
  extra="--no-e2e --steps 5"; [ "$cfg" = cfg2 ] && extra="--no-e2e --no-parity"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2999$N \
      bench.py --gpus $N --config $cfg $extra > gpurun_out/${tag}_bench_${cfg}_n$N.json 2> gpurun_out/${tag}_bench_${cfg}_n$N.err
  echo "bench $cfg rc=$?"; grep -v "^$\|OMP\|\*\*\*" gpurun_out/${tag}_bench_${cfg}_n$N.err | tail -3 | cut -c1-300
  python - <<PY
import json
try:
    d = json.loads([l for l in open('gpurun_out/${tag}_bench_${cfg}_n$N.json') if l.startswith('{')][-1])
    print('$cfg n=$N value %.4g ms/step %.3f parity %s route %s' % (d['value'], d['ms_per_step'], d['parity'].get('checked'), d['exchange']['route']))
except Exception as e:
    print('unreadable:', e)
PY
done
