#!/bin/bash
# weak-scaling bench on N GPUs (usage: gpu_multi_bench.sh N tag)
N=${1:-2}; tag=${2:-r01m}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${tag}_bench_n$N.json 2> gpurun_out/${tag}_bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
for line in open('gpurun_out/${tag}_bench_n$N.json'):
    if line.startswith('{'):
        d=json.loads(line); print(d['n_gpus'], d['value'], d['ms_per_step'], d['roofline']['kernels_ms_per_step_rank0']); print(d['exchange'])
    else: print('STDOUT NOISE:', line[:100])
PY
