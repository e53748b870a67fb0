"""Sampling throughput (BASELINE configs[2]): x+h joint ancestral sampling, T = 1000, 192-atom structures with the
reference's fully connected edge list (parts/train_per_iretation.py:308-313), EGNN L=6 d=128, batch-sharded over the
ranks with no data-path collective.  Prints ONE JSON line: sampled structures/s (whole job) and denoiser graph-steps/s.

    python bench_sampling.py [--structures 64] [--T 1000] [--precision bf16x3]
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench_sampling.py --gpus N

This is the second half of BASELINE.json's metric; bench.py (training step, configs[1]) is the contract line.  The
structure count per GPU is a flag (default 64 so that the run takes seconds; configs[2] is 512 per GPU -- throughput
per structure is flat once the 148 SMs are filled, see profiles/).  Weights: seeded random init, tamed
(oracle-independent copy of the scaling in tests) so that untrained trajectories stay finite.
"""
from __future__ import annotations

import argparse
import json
import os
import time

import torch


def tame_(net, L, fx=0.05, fd=0.02):
    """Scale the coordinate head and the distance columns so an UNTRAINED net gives bounded trajectories (no trained
    weights ship with the reference); code path and shapes are unchanged."""
    with torch.no_grad():
        for l in range(L):
            e = net.egcl_list[l]
            e.mlp_x[4].weight.mul_(fx)
            e.mlp_x[4].bias.mul_(fx)
            e.mlp_m[0].weight[:, -1].mul_(fd)
            e.mlp_x[0].weight[:, -1].mul_(fd)
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import bump_param_epoch
    bump_param_epoch()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--structures", type=int, default=64, help="structures per GPU")
    ap.add_argument("--atoms", type=int, default=192)
    ap.add_argument("--T", type=int, default=1000)
    ap.add_argument("--L", type=int, default=6)
    ap.add_argument("--hidden", type=int, default=128)
    ap.add_argument("--precision", default="bf16x3", choices=["fp32", "bf16x3", "bf16"])
    args = ap.parse_args()

    import torch.distributed as dist
    from diffusion_model_b200 import _lib
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.distributed import init_from_env
    from diffusion_model_b200.sampler import sample_structures

    rank, world, local = init_from_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench_sampling.py: no CUDA device; the product has no CPU path")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    torch.manual_seed(2024)
    H, d, L = 3, args.hidden, args.L
    K0 = 2 * H + 1
    net = EquivariantGNN(L, K0, d, d, K0, d, 1, H + d, d, H).to(dev)
    tame_(net, L)
    net.precision = args.precision
    dp = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=args.T)
    S, n = args.structures, args.atoms
    gen = torch.Generator(device=dev)
    gen.manual_seed(2024 + rank)

    sample_structures(net, dp, n, min(S, 4), generator=gen, device=dev)  # warm-up: builds, packs, allocates
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    l0 = lib.egnn_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    out = sample_structures(net, dp, n, S, generator=gen, device=dev)
    e1.record()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - t0
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    ok = out["finite"].sum().to(torch.float32).view(1)
    if world > 1:
        dist.barrier()
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(ok, op=dist.ReduceOp.SUM)
    sec = float(ms.item()) * 1e-3
    E1 = n * (n - 1)
    K = 2 * H + 1
    flops_step = S * (L * (E1 * (4 * K * d + 4 * d * d + 4 * d) + n * (2 * d * d + 4 * H * d)))
    if rank == 0:
        print(json.dumps(dict(
            metric="sampled structures/s (x+h ancestral sampling, T+1 denoiser calls per structure)",
            value=S * world / sec, unit="structures/s", n_gpus=world, higher_is_better=True, scaling="weak",
            accepted_structures=int(ok.item()), seconds=sec, host_wall_seconds=wall,
            denoiser_graph_steps_per_s=S * world * (args.T + 1) / sec,
            denoiser_tflops=flops_step * (args.T + 1) / sec / 1e12, dtype=args.precision, data="synthetic",
            gpu_launches=int(lib.egnn_launch_count() - l0),
            config=dict(workload="cfg3: x+h joint diffusion sampling, fully connected %d-atom structures, EGNN L=%d d=%d H=3, "
                                 "T=%d, batch-sharded, no collective" % (n, L, d, args.T),
                        structures_per_gpu=S, atoms=n, edges_per_structure=E1, T=args.T, precision=args.precision,
                        weights="seeded random init, coordinate head x0.05 / distance columns x0.02 (untrained nets diverge)"))),
              flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
