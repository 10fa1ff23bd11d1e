#!/bin/bash
# full ncu capture of the scoring variant of the fused kernel (the one the bench times):
# the first fused_kernel launch of the bench is the warm-up step's SCORE variant
tag=${1:-r01b}
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu"
ncu --set full --clock-control none --import-source on -k "regex:fused_kernel|stream_kernel" -s 0 -c 1 -o gpurun_out/${tag}_k12s $CMD > gpurun_out/${tag}_ncu_k12s.log 2>&1
echo "k12 score rc=$?"
ls -la gpurun_out/
