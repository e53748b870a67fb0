"""World-size-2 `gloo` tests (CPU) of the host-side multi-process logic in diffusion_model_b200/distributed.py:
graph sharding, the single flat-gradient all-reduce of a training step, and the parameter broadcast.  The kernels are
not involved (no GPU here); gradients are injected so only the plumbing is exercised."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import egnn_ref


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from diffusion_model_b200 import distributed as D
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.optim import flat_grad_of
    r, w, _ = D.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    try:
        # --- sharding: contiguous, balanced, covers [0, total)
        for total in (0, 1, 7, 64, 65):
            spans = [D.shard_range(total, k, world) for k in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1
        # --- parameter broadcast: ranks start from different weights, end on rank 0's
        torch.manual_seed(100 + rank)
        net = EquivariantGNN(**egnn_ref.net_sizes(2, 32))
        ref0 = [torch.zeros_like(p) for p in net.parameters()]
        if rank == 0:
            ref0 = [p.detach().clone() for p in net.parameters()]
        for t in ref0:
            dist.broadcast(t, src=0)
        D.broadcast_parameters(net, src=0)
        assert all(torch.equal(p.detach(), t) for p, t in zip(net.parameters(), ref0))
        # --- one all-reduce of the flat gradient: per-parameter grads injected as (rank+1) * pattern
        gen = torch.Generator().manual_seed(7)
        pattern = [torch.randn(p.shape, generator=gen) for p in net.parameters()]
        for p, g in zip(net.parameters(), pattern):
            p.grad = (rank + 1) * g.clone()
        flat = D.allreduce_gradients(net)
        scale = sum(k + 1 for k in range(world))
        off = 0
        for p, g in zip(net.parameters(), pattern):
            n = p.numel()
            assert torch.allclose(flat[off:off + n].view(p.shape), scale * g, rtol=1e-6, atol=1e-6)
            assert torch.allclose(p.grad, scale * g, rtol=1e-6, atol=1e-6)   # scattered back into .grad
            off += n
        assert off == flat.numel() == net.flat_parameters().numel()
        g2, _ = flat_grad_of(net)
        assert torch.equal(g2, flat)
        # --- flat buffer all-reduce helper is a no-op contract when called on a 1-element group of tensors
        one = torch.full((3,), float(rank + 1))
        D.allreduce_flat(one)
        assert torch.equal(one, torch.full((3,), float(scale)))
        out.put((rank, "ok"))
    except Exception as e:  # surface the failure in the parent
        out.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_sharding_broadcast_and_gradient_allreduce():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = dict(out.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert results == {0: "ok", 1: "ok"}, results
