#!/bin/bash
tag=${1:-r02i}
mkdir -p gpurun_out
bash tools/gpu_variants.sh $tag 0 - _x
unset LIME_B200_LIB
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "xchg or deferred" > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest.log
timeout 300 python tools/k12_modes.py > gpurun_out/${tag}_modes.log 2>&1; echo "modes rc=$?"; cat gpurun_out/${tag}_modes.log | cut -c1-420
timeout 300 ncu --set full --clock-control none --import-source on -k "regex:stream_kernel<\(int\)3" -s 2 -c 1 -o gpurun_out/${tag}_k12flat python tools/k12_modes.py > gpurun_out/${tag}_ncu_flat.log 2>&1; echo "flat capture rc=$?"
