#!/bin/bash
# Round 2, one GPU: ncu --set full of the exchange kernels (route / apply) and of the record-appending stream kernel,
# in loopback with 8 virtual owners (tools/k12_modes.py)
tag=${1:-r02af}
mkdir -p gpurun_out
timeout 300 python tools/k12_modes.py > gpurun_out/${tag}_modes.log 2>&1; echo "modes rc=$?"; cut -c1-300 gpurun_out/${tag}_modes.log | tail -8
# world = 1 comes first (14 route launches, 98 apply launches): skip into the world = 8 part
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:route_kernel" -s 16 -c 1 -o gpurun_out/${tag}_xroute python tools/k12_modes.py > gpurun_out/${tag}_ncu_xroute.log 2>&1; echo "route capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:apply_kernel" -s 100 -c 1 -o gpurun_out/${tag}_xapply python tools/k12_modes.py > gpurun_out/${tag}_ncu_xapply.log 2>&1; echo "apply capture rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:stream_kernel<\(int\)3" -s 16 -c 1 -o gpurun_out/${tag}_k12flat python tools/k12_modes.py > gpurun_out/${tag}_ncu_flat.log 2>&1; echo "flat capture rc=$?"
