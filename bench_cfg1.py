#!/usr/bin/env python
"""BASELINE.json configs[0] -- the reference's own CPU-runnable case: EGNN 4 layers hidden 64, fully connected graphs of 24
atoms, batch 8, the full T = 1000 ancestral trajectory (1001 denoiser calls, parts/train_per_iretation.py:299-428).

Times, in the same run on the same box:
  * the CPU restatement of the reference path (oracle/, all host threads; BASELINE.md 3 names this trajectory as the
    headline CPU number), with the normaliser of EGCL.message per structure (8 x batch-1 = what generate() computes) and
    over the whole call (a batch of 8 in one reference call);
  * the CUDA sampler on the same 8 structures, both normaliser scopes, and on 512 / 4096 structures (what one GPU can do
    with the same 1001 launches-bound steps).
Prints ONE JSON line.  This is a parity-config timing for profiles/, not the contract bench line (bench.py).

    python bench_cfg1.py [--skip-cpu] [--precision bf16x3]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--precision", default="bf16x3", choices=["fp32", "bf16x3", "bf16"])
    ap.add_argument("--T", type=int, default=1000)
    args = ap.parse_args()
    import bench
    from diffusion_model_b200 import _lib
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.sampler import sample_structures
    if not torch.cuda.is_available():
        raise SystemExit("bench_cfg1.py: no CUDA device; the product has no CPU path")
    dev = torch.device("cuda", 0)
    lib = _lib.load()
    L, d, H, n, S, T = 4, 64, 3, 24, 8, args.T
    torch.manual_seed(2024)
    net = EquivariantGNN(**bench.net_sizes(L, d, H)).to(dev)
    bench.tame_(net, L)
    net.precision = args.precision
    sd = {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    dp = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=T)
    out = dict(config=dict(workload=f"cfg1: EGNN L={L} d={d} H=3, {S} x {n}-atom fully connected graphs (E = {S * n * (n - 1)}), "
                                    f"T = {T} ancestral sampling (T+1 denoiser calls)", precision=args.precision), gpu={}, cpu={})
    for scope in ("graph", "call"):
        for structures in ((S, 512, 4096) if scope == "graph" else (S,)):
            gen = torch.Generator(device=dev)
            gen.manual_seed(7)
            sample_structures(net, dp, n, structures, generator=gen, device=dev, norm_scope=scope)  # warm-up + capture
            torch.cuda.synchronize(dev)
            l0 = lib.egnn_launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            r = sample_structures(net, dp, n, structures, generator=gen, device=dev, norm_scope=scope)
            e1.record()
            torch.cuda.synchronize(dev)
            sec = e0.elapsed_time(e1) * 1e-3
            out["gpu"][f"{scope}_S{structures}"] = dict(
                structures_per_s=structures / sec, denoiser_graph_steps_per_s=structures * (T + 1) / sec, seconds=sec,
                host_wall_seconds=time.perf_counter() - t0, us_per_sampler_step=1e6 * sec / (T + 1),
                accepted=int(r["finite"].sum()), launches_per_step=(lib.egnn_launch_count() - l0) / (T + 1))
    if not args.skip_cpu:
        from oracle import diffusion_ref as D, loops_ref as LP
        info = bench.host_info()
        torch.set_num_threads(info["logical_cpus"])
        sch = D.Schedule(0.05, 2.0, T)
        gen = torch.Generator().manual_seed(11)
        per = [LP.make_noises(n, T, 2, gen) for _ in range(S)]
        cat = {k: torch.cat([p[k] for p in per], dim=1 if k.startswith("step") else 0) for k in per[0]}
        for scope in ("graph", "call"):
            t0 = time.perf_counter()
            with torch.no_grad():
                _, _, fin = LP.sample_structures_batched(sd, L, sch, n, S, cat, norm_scope=scope)
            sec = time.perf_counter() - t0
            out["cpu"][f"{scope}_S{S}"] = dict(structures_per_s=S / sec, denoiser_graph_steps_per_s=S * (T + 1) / sec, seconds=sec,
                                               accepted=int(fin.sum()), cores=info["logical_cpus"], kind="port",
                                               cpu_model=info["cpu_model"], physical_cores=info["physical_cores"])
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
