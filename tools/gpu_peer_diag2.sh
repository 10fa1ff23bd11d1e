#!/bin/bash
# where does the time of the peer-memory adds go: the real kernel with the adds re-routed (LIME_DEBUG)
N=${1:-8}; tag=${2:-r01s}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
for dbg in ${3:-16 64 128 256}; do
  LIME_DEBUG=$dbg LIME_NO_DIAG=1 timeout 200 $TR bench.py --gpus $N --steps 2 --warmup 1 > gpurun_out/${tag}_bench_n${N}_dbg$dbg.json 2> gpurun_out/${tag}_bench_n${N}_dbg$dbg.err; echo "bench dbg=$dbg rc=$?"
  python - <<PY
import json
for line in open('gpurun_out/${tag}_bench_n${N}_dbg$dbg.json'):
    if line.startswith('{'):
        d=json.loads(line); print($dbg, d['n_gpus'], d['ms_per_step'], d['roofline']['k12_ms_per_launch_by_rank'])
PY
done
