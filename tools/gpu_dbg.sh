#!/bin/bash
# experiment: time the fused kernel with parts switched off (LIME_DEBUG bits)
mkdir -p gpurun_out
for d in "$@"; do
LIME_DEBUG=$d python - <<PY
import torch, numpy as np, os, sys
sys.path.insert(0, '.')
from lime_b200 import datagen
from lime_b200.api import LimeContext, MEM_DEVICE
n, nr, ng = 402_000_000, 2_000_000, 100
e,l,d = datagen.synth_arrays(n, nr, ng, 100, seed=20261106, device='cuda')
ctx = LimeContext(0); st = torch.cuda.current_stream(); ctx.set_stream(st.cuda_stream)
ts=[]
for it in range(4):
    ctx.load(l.data_ptr(), d.data_ptr(), e.data_ptr(), nr, ng, mem=MEM_DEVICE, n_total=n)
    a,b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record(st)
    try:
        r = ctx.cluster_similarity(16, 1, False)
    except Exception as ex:
        r = str(ex)[:60]
    b.record(st); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
print('LIME_DEBUG=%s' % os.environ.get('LIME_DEBUG'), 'ms', [round(t,3) for t in ts], r, ctx.sim_stats())
PY
done
