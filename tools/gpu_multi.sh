#!/bin/bash
# N-GPU run: parity check of the sharded job, then the weak-scaling bench (usage: gpu_multi.sh N tag)
N=${1:-2}; tag=${2:-r01m}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR tools/gpu_multi_check.py > gpurun_out/${tag}_check_n$N.log 2>&1; echo "check rc=$?"; grep -E "rank|parity|Error|error" gpurun_out/${tag}_check_n$N.log | tail -8
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${tag}_bench_n$N.json 2> gpurun_out/${tag}_bench_n$N.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench_n$N.err
python - <<PY
import json
try:
    d=json.load(open('gpurun_out/${tag}_bench_n$N.json')); print(d['n_gpus'], d['value'], d['ms_per_step'], d['roofline']['kernels_ms_per_step_rank0'], d['exchange'])
except Exception as e: print('bench parse failed', e)
PY
