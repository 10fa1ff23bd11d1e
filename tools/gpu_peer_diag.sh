#!/bin/bash
# diagnosis of the peer-memory job on N GPUs: add-rate probe, bench, bench without the adds
N=${1:-4}; tag=${2:-r01r}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 200 $TR tools/gpu_peer_probe.py 2>&1 | grep -E "^\[" > gpurun_out/${tag}_peer_probe_n$N.log; tail -12 gpurun_out/${tag}_peer_probe_n$N.log
for dbg in 0 2; do
  LIME_DEBUG=$dbg timeout 300 $TR bench.py --gpus $N --steps 3 --warmup 2 > gpurun_out/${tag}_bench_n${N}_dbg$dbg.json 2> gpurun_out/${tag}_bench_n${N}_dbg$dbg.err; echo "bench dbg=$dbg rc=$?"
  python - <<PY
import json
for line in open('gpurun_out/${tag}_bench_n${N}_dbg$dbg.json'):
    if line.startswith('{'):
        d=json.loads(line); print(d['n_gpus'], d['value'], d['ms_per_step'], d['roofline']['k12_ms_per_launch_by_rank']); print(d['exchange']['phases_ms_max_over_ranks'])
PY
done
