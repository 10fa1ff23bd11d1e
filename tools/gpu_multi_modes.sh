#!/bin/bash
# weak-scaling bench on N GPUs in both multi-GPU modes (usage: gpu_multi_modes.sh N tag)
N=${1:-4}; tag=${2:-r01t}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for mode in ${3:-exchange peer}; do
  LIME_MULTI=$mode timeout 300 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${tag}_bench_n${N}_$mode.json 2> gpurun_out/${tag}_bench_n${N}_$mode.err; echo "bench $mode rc=$?"
  python - <<PY
import json
for line in open('gpurun_out/${tag}_bench_n${N}_$mode.json'):
    if line.startswith('{'):
        d=json.loads(line); print('$mode', d['n_gpus'], d['value'], d['ms_per_step'], d['roofline']['kernels_ms_per_step_rank0']); print(d['exchange']['phases_ms_max_over_ranks'])
PY
done
