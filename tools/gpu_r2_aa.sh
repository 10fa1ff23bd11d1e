#!/bin/bash
# Round 2: N-GPU default line with its e2e leg (after the e2e leg went back to one job at a time)
tag=${1:-r02aa}; N=${2:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29991 \
    bench.py --gpus $N > gpurun_out/${tag}_bench_cfg2_n$N.json 2> gpurun_out/${tag}_bench_cfg2_n$N.err
echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench_cfg2_n$N.err | cut -c1-300
python - <<PY
import json
d = json.loads([l for l in open('gpurun_out/${tag}_bench_cfg2_n$N.json') if l.startswith('{')][-1])
print('n=$N value %.4g ms/step %.3f parity %s e2e %s' % (d['value'], d['ms_per_step'], d['parity'].get('checked'), json.dumps(d.get('e2e'))[:400]))
PY
