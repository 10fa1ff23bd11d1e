#!/bin/bash
# Round profile: plain bench run, ncu launch list of the same command, full capture of the top kernels.
# usage (under gpurun): bash tools/gpu_profile.sh <tag>
tag=${1:-r01}
mkdir -p gpurun_out
K="regex:stream_kernel|bytesum_kernel|fused_kernel|warp_kernel|block_kernel|count_kernel|emit_kernel|choose_kernel|choose_sparse_kernel|classify_kernel|narrow_kernel"
CMD="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 200 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k "regex:stream_kernel" -s 0 -c 1 -o gpurun_out/${tag}_k12 $CMD > gpurun_out/${tag}_ncu_k12.log 2>&1
echo "k12 rc=$?"
ncu --set full --clock-control none --import-source on -k "regex:choose_sparse_kernel|choose_kernel" -s 1 -c 1 -o gpurun_out/${tag}_k3 $CMD > gpurun_out/${tag}_ncu_k3.log 2>&1
echo "k3 rc=$?"
