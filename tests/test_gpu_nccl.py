"""NCCL-level correctness: 2 ranks, one per GPU, over NVLink.  Skipped on a one-GPU box; the host-side
sharding logic is covered on CPU by tests/test_distributed_cpu.py (gloo, world size 2).  The checks themselves live in
tests/nccl_worker.py: DDP-step gradients == the single-GPU gradients of the whole minibatch (norm_scope='graph'), and
batch-sharded sampling == the unsharded run under injected noise."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (NCCL)")
@pytest.mark.parametrize("precision", ["bf16x3", "fp32"])
def test_ddp_gradients_and_sharded_sampler_over_nccl(lib, precision):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "nccl_worker.py"), precision]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    print(r.stdout[-2000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "nccl ok world=2" in r.stdout
