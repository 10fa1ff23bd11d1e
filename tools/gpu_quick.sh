#!/bin/bash
# quick GPU check: parity tests + a short resident-input bench (run under gpurun)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu "$@" > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_q.err
python -c "
import json; d=json.load(open('gpurun_out/bench_q.json')); print(d['value'], d['ms_per_step'], d['roofline']['kernels_ms_per_step'], d['roofline']['frac'], d['gpu_launches'], d['roofline']['kernel_time_fraction_of_step'])"
