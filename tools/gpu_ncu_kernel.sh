#!/bin/bash
# full ncu capture of one kernel of the bench: gpu_ncu_kernel.sh <tag> <kernel-regex> [skip]
tag=$1; k=$2; skip=${3:-0}
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu"
ncu --set full --clock-control none --import-source on -k "regex:$k" -s $skip -c 1 -o gpurun_out/${tag} $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu $k rc=$?"
