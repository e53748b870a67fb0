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


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_second_gpu_of_one_process(lib):
    """one process driving a GPU that is not the current device.  The ctypes layer makes the tensors' device current
    for the call (kernel attributes and the static scratch address are per device inside the library): forward, backward and
    the allocation-free graph preparation on cuda:1 while cuda:0 stays current, equal to the same work on cuda:0."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200._lib import ptr, stream_ptr, call
    from diffusion_model_b200.graph import build_edges_fc
    from oracle import egnn_ref
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    sizes = egnn_ref.net_sizes(2, 128)
    sd = egnn_ref.init_state_dict(**sizes)
    gen = torch.Generator().manual_seed(1)
    h0, x0 = torch.randn(60, 3, generator=gen), torch.randn(60, 3, generator=gen)
    res = {}
    for dev in ("cuda:0", "cuda:1"):
        net = EquivariantGNN(**sizes)
        net.load_state_dict(sd)
        net = net.to(dev)
        ei = build_edges_fc([20, 40], dev)
        h, x = h0.to(dev).requires_grad_(True), x0.to(dev).requires_grad_(True)
        ho, xo = net(ei, h, x)
        (ho.square().sum() + xo.square().sum()).backward()
        assert torch.cuda.current_device() == 0
        E = ei.shape[1]
        o = [torch.empty(E, dtype=torch.int32, device=dev) for _ in range(2)] + \
            [torch.empty(61, dtype=torch.int32, device=dev) for _ in range(2)] + \
            [torch.empty(E, dtype=torch.int32, device=dev), torch.empty(2, dtype=torch.int32, device=dev)]
        call("egnn_graph_prepare_noalloc", ptr(ei), 60, E, *[ptr(t) for t in o], stream_ptr(torch.device(dev)))
        torch.cuda.synchronize(dev)
        res[dev] = [ho.detach().cpu(), xo.detach().cpu(), h.grad.cpu(), net.flat_parameters().grad.cpu() if net.flat_parameters().grad is not None
                    else torch.cat([p.grad.reshape(-1) for p in net.parameters()]).cpu()] + [t.cpu() for t in o]
    for a, b in zip(res["cuda:0"], res["cuda:1"]):
        assert torch.equal(a, b)
