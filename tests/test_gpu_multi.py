"""GPU, >= 2 devices: the position-sharded job under torchrun (one rank per GPU, NCCL over NVLink) against the
single-GPU job on the whole arrays -- assignments, nClusters and maxLen bit for bit, for both routes (the record
exchange XchgJob with CUDA-IPC mapped inboxes, and the exact triples route ShardedJob).  The work is done by
tools/gpu_multi_check.py; this test launches it with min(device_count, 8) ranks and skips itself below 2."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(1500)]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_sharded_job_equals_single_gpu_job_under_torchrun():
    import torch
    n = min(torch.cuda.device_count(), 8)
    if n < 2:
        pytest.skip("needs at least 2 CUDA devices")
    env = dict(os.environ)
    env.pop("LIME_MULTI", None)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
                        os.path.join(ROOT, "tools", "gpu_multi_check.py")], cwd=ROOT, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1400)
    assert r.returncode == 0, r.stdout[-6000:]
    assert "multi-GPU parity ok" in r.stdout, r.stdout[-6000:]
