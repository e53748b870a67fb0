"""Worker of tests/test_gpu_nccl.py, launched as `python -m torch.distributed.run --nproc-per-node 2 tests/nccl_worker.py`
(one rank per GPU, NCCL).  Exits non-zero on any mismatch; rank 0 prints one summary line.

  1. data-parallel step: a global minibatch of 2W graphs, two per rank; loss = SSE / (graphs of the rank)
     (parts/train_per_iretation.py:169), flat-gradient all-reduce (sum), scaled by 1/W  ==  the gradient of ONE GPU
     running the whole minibatch, with the normaliser of EGCL.message per graph (norm_scope='graph': shard-independent).
  2. batch-sharded sampling with injected noise  ==  the unsharded run of the same structures on one GPU.
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def main():
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.data import Batch, Data
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.distributed import (allreduce_gradients, broadcast_parameters, init_from_env, sample_sharded,
                                                  shard_range)
    from diffusion_model_b200.graph import build_edges_fc
    from diffusion_model_b200.optim import FlatAdam, flat_grad_of
    from diffusion_model_b200.sampler import sample_structures
    from diffusion_model_b200.train_per_iretation import batch_loss
    import bench
    rank, world, local = init_from_env("nccl")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    precision = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
    L, d, H, A = 3, 64, 3, 2
    torch.manual_seed(100 + rank)  # different init per rank on purpose: broadcast_parameters must fix it
    net = EquivariantGNN(**bench.net_sizes(L, d, H)).to(dev)
    net.precision = precision
    broadcast_parameters(net)
    bench.tame_(net, L)
    w = net.flat_parameters().clone()
    dist.all_reduce(w, op=dist.ReduceOp.MAX)
    assert torch.equal(w, net.flat_parameters()), "parameters differ across ranks after broadcast"

    # ---- 1. DDP step ------------------------------------------------------------------------------------------
    T = 1000
    dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=T)
    gen = torch.Generator().manual_seed(5)  # same on every rank: the global minibatch and its draws
    G = 2 * world
    graphs, t_all, npos, nh = [], [], [], []
    for g in range(G):
        n = 20 + 3 * g
        x = torch.nn.functional.one_hot(torch.randint(0, 2, (n,), generator=gen), 2).float()
        graphs.append(Data(x=x, pos=4.0 * torch.rand(n, 3, generator=gen), edge_index=build_edges_fc([n], dev).cpu()))
        t_all.append(int(torch.randint(1, T + 1, (1,), generator=gen)))
        npos.append(torch.randn(n, 3, generator=gen))
        nh.append(torch.randn(n, A, generator=gen))
    prm = dict(atom_type_size=A, norm_scope="graph")

    def grad_of(idx):
        b = Batch.from_data_list([graphs[i] for i in idx]).to(dev)
        net.zero_grad(set_to_none=True)
        net._last_grad_flat = None
        loss, _, _ = batch_loss(net, b, dp, prm, t=[t_all[i] for i in idx], noise_pos=torch.cat([npos[i] for i in idx]).to(dev),
                                noise_h=torch.cat([nh[i] for i in idx]).to(dev))
        loss.backward()
        return loss.detach()

    lo, hi = shard_range(G, rank, world)
    loss_r = grad_of(range(lo, hi))
    g_ddp = allreduce_gradients(net).clone() / world
    loss_sum = loss_r.clone()
    dist.all_reduce(loss_sum)
    loss_full = grad_of(range(G))           # every rank also runs the whole minibatch alone
    g_full = flat_grad_of(net)[0].clone()
    e_g, e_l = rel(g_ddp, g_full), abs(float(loss_sum) / world - float(loss_full)) / abs(float(loss_full))
    assert e_g <= 2e-5 and e_l <= 1e-5, ("ddp gradient / loss", e_g, e_l)
    # the fused optimizer step is deterministic: identical gradients on every rank keep the replicas bit-identical
    opt = FlatAdam(net, lr=1e-3)
    opt.step(grad_scale=1.0 / world)
    w = net.flat_parameters().clone()
    wmax = w.clone()
    dist.all_reduce(wmax, op=dist.ReduceOp.MAX)
    assert torch.equal(w, wmax), "ranks diverged after the optimizer step"

    # ---- 2. sharded sampling under injected noise ------------------------------------------------------------------
    from oracle import loops_ref as LP  # make_noises only (the checker's draw layout); nothing of the product uses it
    Ts, n, S = 25, 16, 6
    dps = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=Ts)
    gen = torch.Generator().manual_seed(9)
    per = [LP.make_noises(n, Ts, A, gen) for _ in range(S)]
    cat = {k: torch.cat([p[k] for p in per], dim=1 if k.startswith("step") else 0) for k in per[0]}
    with torch.no_grad():
        out = sample_sharded(net, dps, n, S, noises=cat, device=dev)
        full = sample_structures(net, dps, n, S, noises=cat, device=dev)
    assert out["range"] == shard_range(S, rank, world)
    e_p = rel(out["pos_all"], full["pos"])
    assert e_p <= 1e-4, ("sharded sampler", e_p)
    assert torch.equal(out["atoms_all"], full["atoms"]) and bool(out["finite_all"].all())
    lo, hi = out["range"]
    assert torch.equal(out["pos_all"][lo:hi], out["pos"])
    if rank == 0:
        print(f"nccl ok world={world} {precision}: ddp grad {e_g:.1e} loss {e_l:.1e} sharded sampler {e_p:.1e}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
