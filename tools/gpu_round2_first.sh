#!/bin/bash
# First GPU call of the next round: everything that was written after round 1's GPU minutes were spent.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round2_first.sh'
# 1. the whole GPU suite, 2. the narrowed lcp transfer with its xfail lifted, 3. the bench line (both e2e
# modes under e2e.modes), 4. the drop-in executables with LIME_NARROW_LCP=1 against their plain outputs,
# 5. the compile-time experiments of the stream kernel.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 gpurun_out/r2_pytest_gpu.log
python -m pytest tests/test_zz_gpu_narrow_load.py -m gpu -q --runxfail > gpurun_out/r2_pytest_narrow.log 2>&1; echo "narrow rc=$?"; tail -5 gpurun_out/r2_pytest_narrow.log
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d = json.load(open('gpurun_out/r2_bench_n1.json'))
    print('value %.4g  e2e %.4g  modes %s' % (d['value'], d['e2e']['value'], json.dumps(d['e2e'].get('modes'))))
except Exception as e:
    print('bench line unreadable:', e)
PY
LIME_NARROW_LCP=1 python -m pytest tests/test_host_cli.py -m gpu -x -q > gpurun_out/r2_cli_narrow.log 2>&1; echo "cli (LIME_NARROW_LCP=1) rc=$?"; tail -2 gpurun_out/r2_cli_narrow.log
bash tools/gpu_exp.sh 1 2 4 8 12 15
