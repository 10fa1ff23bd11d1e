#!/bin/bash
# what the driver runs at round end: default bench (ours), reference arm; then the round profile
tag=${1:-r01h}
mkdir -p gpurun_out
( time python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err ) 2> gpurun_out/${tag}_bench.time; echo "bench rc=$?"; tail -2 gpurun_out/${tag}_bench.err; cat gpurun_out/${tag}_bench.time | grep real
( time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err ) 2> gpurun_out/${tag}_ref.time; echo "ref rc=$?"; tail -2 gpurun_out/${tag}_ref.err; cat gpurun_out/${tag}_ref.time | grep real
bash tools/gpu_profile.sh $tag
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${tag}_smoke.log
