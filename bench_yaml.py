#!/usr/bin/env python
"""Training step of the ONLY configuration the reference ships (parameters.yaml through main.py:102-121): conditional model,
h = [2 atom types | 32 compressed spectrum | exO | t] (H = 36, K0 = 73), L = 5, m_hidden = x_hidden = h_hidden = 1024,
m_size = 256, SpectrumCompressor 200 -> 150 -> 100 -> 50 -> 32, batch_size 1, RAdamScheduleFree, lr 1e-5.

These widths have no fused tcgen05 kernel (2H+1 > 16, Mh != Ms, d = 1024): under `precision = 'auto'` (= bf16x3) the forward
and the backward of the edge MLPs run the tile-GEMM tcgen05 path of csrc/tc_wide.cu; `--precision fp32` runs the CUDA-core
kernels throughout.  Prints ONE JSON line: graph-steps/s of the eps-prediction training step on a
fully connected graph of `--atoms` atoms (the reference's datasets hold clusters of <= ~120 atoms) for the GPU path and for
the CPU oracle on the same step.  A parity-config timing for profiles/, not the contract bench line (bench.py).

    python bench_yaml.py [--atoms 96] [--batch 1] [--steps 10] [--precision auto|fp32|bf16x3|bf16] [--skip-cpu]
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

YAML = dict(L=5, atom_type_size=2, compressed_spectrum_size=32, spectrum_size=200, compressor_hidden_dim=[150, 100, 50], exO_size=1,
            t_size=1, d_size=1, m_size=256, m_hidden_size=1024, h_hidden_size=1024, x_hidden_size=1024, T=1000, s=1e-5, power=2.0, lr=1e-5)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--atoms", type=int, default=96)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--precision", default="auto")
    ap.add_argument("--sample-structures", type=int, default=0,
                    help="also sample this many structures in one batch through the public sampler (T = 1000 + the t = 0 step, "
                         "conditioning held constant): BASELINE.md's published figure is this loop at ONE structure per call")
    args = ap.parse_args(argv)
    import bench
    from diffusion_model_b200 import _lib
    from diffusion_model_b200.DataPreprocessor import SpectrumCompressor
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.data import Batch, Data
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.graph import build_edges_fc
    from diffusion_model_b200.optim import RAdamScheduleFree
    from diffusion_model_b200.train_per_iretation import _conditioning, batch_loss
    if not torch.cuda.is_available():
        raise SystemExit("bench_yaml.py: no CUDA device; the product has no CPU path")
    # data parallel under torchrun (weak scaling: --batch graphs on EVERY rank): the flat EGNN gradient (36 MB) is all-reduced
    # in three buckets of layers on a communication stream while backward runs (distributed.OverlappedGradSync), the
    # compressor's few small gradients in one call behind it
    import torch.distributed as dist
    from diffusion_model_b200.distributed import OverlappedGradSync, broadcast_parameters, init_from_env
    rank, world, local = init_from_env("nccl")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    y = YAML
    H = y["atom_type_size"] + y["compressed_spectrum_size"] + y["t_size"] + y["exO_size"]      # main.py:102-110
    K0 = 2 * H + y["d_size"]
    sizes = dict(L=y["L"], m_input=K0, m_hidden=y["m_hidden_size"], m_output=y["m_size"], x_input=K0, x_hidden=y["x_hidden_size"],
                 x_output=1, h_input=H + y["m_size"], h_hidden=y["h_hidden_size"], h_output=H)
    torch.manual_seed(2024)
    net = EquivariantGNN(**sizes).to(dev)
    comp = SpectrumCompressor(y["spectrum_size"], y["compressor_hidden_dim"], y["compressed_spectrum_size"]).to(dev)
    net.precision = args.precision
    prec = net.resolved_precision()
    sd = {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    cw = {k: v.detach().cpu().clone() for k, v in comp.state_dict().items()}
    params = dict(atom_type_size=2, conditional=True, to_compress_spectrum=True, give_exO=True, optimizer="RAdamScheduleFree")
    dp = E3DiffusionProcess(s=y["s"], power=y["power"], num_diffusion_timestep=y["T"])
    gen = torch.Generator().manual_seed(7 + rank)
    n, G = args.atoms, args.batch
    graphs = []
    for _ in range(G):
        pos, onehot, _ = bench.synth_batch(1, n, int(torch.randint(0, 10 ** 6, (1,), generator=gen)))
        exo = torch.zeros(n, 1)
        exo[0] = 1.0
        graphs.append(Data(x=onehot, pos=pos, edge_index=build_edges_fc([n], dev).cpu(), spectrum=torch.rand(n, y["spectrum_size"], generator=gen),
                           exO=exo))
    batch = Batch.from_data_list(graphs).to(dev)
    E = int(batch.edge_index.shape[1])
    opt = RAdamScheduleFree(list(net.parameters()) + list(comp.parameters()), lr=y["lr"])
    opt.train()
    nn_dict = {"egnn": net, "spectrum_compressor": comp}

    sync = None
    if world > 1:
        broadcast_parameters(net)
        for p_ in comp.parameters():
            dist.broadcast(p_.data, src=0)
        sync = OverlappedGradSync(net, buckets=3)
    comp_params = list(comp.parameters())

    def step():
        opt.zero_grad(set_to_none=True) if world > 1 else opt.zero_grad()
        loss, _, _ = batch_loss(net, batch, dp, params, cond=_conditioning(batch, params, nn_dict))
        (loss / world).backward()  # the all-reduce sums: mean over the global minibatch
        if world > 1:
            flat = torch.cat([p_.grad.reshape(-1) for p_ in comp_params])
            dist.all_reduce(flat)
            off = 0
            for p_ in comp_params:
                p_.grad.copy_(flat[off:off + p_.numel()].view_as(p_))
                off += p_.numel()
            sync.finish()
        opt.step()
        return loss

    for _ in range(3):
        step()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    l0 = lib.egnn_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks = bench.ClockSampler(local)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize(dev)
    clk = clocks.stop()
    ms = e0.elapsed_time(e1) / args.steps
    if world > 1:
        t_ = torch.tensor([ms], device=dev)
        dist.barrier()
        dist.all_reduce(t_, op=dist.ReduceOp.MAX)
        ms = float(t_.item())
    launches = int(lib.egnn_launch_count() - l0)
    # one denoiser call alone (what every sampler step costs): the network forward on the batch's own input
    with torch.no_grad():
        cond = _conditioning(batch, params, nn_dict)
        hin = torch.cat((batch.x, cond, torch.full((G * n, 1), 0.5, device=dev)), dim=1)
        for _ in range(2):
            net(batch.edge_index, hin, batch.pos, batch=batch.batch, norm_scope="graph")
        e0.record()
        for _ in range(args.steps):
            net(batch.edge_index, hin, batch.pos, batch=batch.batch, norm_scope="graph")
        e1.record()
        torch.cuda.synchronize(dev)
    fwd_ms = e0.elapsed_time(e1) / args.steps
    P = sum(p.numel() for p in net.parameters())
    K = K0
    flops_fwd = y["L"] * (E * (2 * K * 1024 + 2 * 1024 * 256 + 2 * 256 + 2 * K * 1024 + 2 * 1024 * 1024 + 2 * 1024)
                          + G * n * (2 * (H + 256) * 1024 + 2 * 1024 * H))          # SURVEY.md 8(d), general widths
    peaks = bench.measured_peaks()
    tf_train, tf_fwd = 3 * flops_fwd / (ms * 1e-3) / 1e12, flops_fwd / (fwd_ms * 1e-3) / 1e12
    # tensor-bound path: algorithmic FLOPs (general widths) over the measured time of ALL kernels
    # of the step / of one denoiser call, against the measured bf16 peak.  bf16x3 issues 3 MMA products per algorithmic one
    # (and the backward recomputes the forward GEMMs): `issued_frac` counts what the tensor pipe actually executes.
    roofline = dict(bound="tensor", kernel="tc_wide_gemm (all epilogue variants; whole training step incl. gathers, column sums, optimizer)",
                    achieved=tf_train, peak=peaks["tflops"], unit="TFLOP/s", frac=tf_train / peaks["tflops"], traffic=None,
                    peak_source=f"{peaks['source']} bf16 sustained",
                    issued_frac=(3 if prec == "bf16x3" else 1) * (4.0 / 3.0) * tf_train / peaks["tflops"] if prec != "fp32" else None,
                    forward=dict(achieved=tf_fwd, frac=tf_fwd / peaks["tflops"],
                                 issued_frac=(3 if prec == "bf16x3" else 1) * tf_fwd / peaks["tflops"] if prec != "fp32" else None),
                    ncu="profiles/r2_ncu_wide_gemm.csv: the [E x 1024] x [1024 x 1024] GEMM runs at 83.6 % tensor-pipe active")
    out = dict(metric="denoiser graph-steps/s (training fwd+bwd+optimizer), parameters.yaml network", higher_is_better=True,
               steps=args.steps, warmup=3, data="synthetic", vs_baseline=None, clocks=clk, roofline=roofline, value=G * world / (ms * 1e-3),
               unit="graph-steps/s", n_gpus=world, scaling="weak", ms_per_step=ms,
               dtype="fp32 (CUDA-core FFMA kernels)" if prec == "fp32" else f"{prec} (tcgen05 tile GEMMs, csrc/tc_wide.cu: edge and node MLPs, forward and backward)",
               gpu_launches=launches, loss=float(loss.detach()), parameters=P, train_tflops_per_gpu=3 * flops_fwd / (ms * 1e-3) / 1e12,
               denoiser_forward_ms=fwd_ms, forward_tflops=flops_fwd / (fwd_ms * 1e-3) / 1e12,
               config=dict(workload=f"parameters.yaml: conditional EGNN L=5, K0=73, m/x/h hidden 1024, m_size 256, SpectrumCompressor, "
                                    f"RAdamScheduleFree; {G} fully connected graph(s) of {n} atoms (E = {E}), T=1000", atoms=n, batch=G, edges=E))
    if args.sample_structures > 0 and world == 1:
        # generate()'s loop (parts/train_per_iretation.py:299-442) batched: the reference's published run is this, one structure
        # per call (~6.9 ms per denoiser graph-step, ~0.144 structures/s on the author's GPU, hidden sizes not recorded)
        import copy
        from diffusion_model_b200.sampler import sample_structures
        S = args.sample_structures
        snet = copy.deepcopy(net)
        bench.tame_(snet, y["L"], fx=0.002, fd=0.001)
        snet.precision = args.precision
        with torch.no_grad():
            cond1 = _conditioning(batch, params, nn_dict)[:n]
        gen_d = torch.Generator(device=dev)
        gen_d.manual_seed(11)
        l1 = lib.egnn_launch_count()
        e0.record()
        res = sample_structures(snet, dp, n, S, atom_type_size=2, cond=cond1, generator=gen_d, device=dev)
        e1.record()
        torch.cuda.synchronize(dev)
        sec_s = e0.elapsed_time(e1) * 1e-3
        T1 = y["T"] + 1
        out["sampling"] = dict(metric="sampled structures/s (conditional x+h ancestral sampling, T+1 denoiser calls per structure)",
                               value=S / sec_s, unit="structures/s", seconds=sec_s, structures=S, accepted=int(res["finite"].sum()),
                               denoiser_graph_steps_per_s=S * T1 / sec_s, ms_per_denoiser_call=1e3 * sec_s / T1,
                               gpu_launches=int(lib.egnn_launch_count() - l1),
                               published_reference="~144 denoiser graph-steps/s, ~0.144 structures/s, one structure per call (BASELINE.md, author's GPU, "
                                                   "hidden sizes of that run not recorded)",
                               weights="seeded random init, coordinate head x0.002 / distance columns x0.001 (untrained nets diverge)")
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if not args.skip_cpu and world == 1:
        from oracle import diffusion_ref as D, egnn_ref, loops_ref as LP
        from oracle.radam_sf_ref import RAdamScheduleFreeRef
        info = bench.host_info()
        torch.set_num_threads(info["logical_cpus"])
        p = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        cwp = {k: v.clone().requires_grad_(True) for k, v in cw.items()}
        ro = RAdamScheduleFreeRef(list(p.values()) + list(cwp.values()), lr=y["lr"])
        ro.train()
        sch = D.Schedule(y["s"], y["power"], y["T"])
        b = Batch.from_data_list(graphs)
        idx = sorted({int(k.split(".")[1]) for k in cwp})

        def compress(spec):
            z = spec
            for q, i in enumerate(idx):
                z = z @ cwp[f"mlp.{i}.weight"].t() + cwp[f"mlp.{i}.bias"]
                if q + 1 < len(idx):
                    z = torch.relu(z)
            return z
        g2 = torch.Generator().manual_seed(1)
        times = []
        for k in range(3):
            t0 = time.perf_counter()
            ro.zero_grad()
            cond = torch.cat((compress(b.spectrum), b.exO), dim=1)
            t = torch.randint(1, y["T"] + 1, (G,), generator=g2)
            loss_c, _ = LP.train_loss(lambda e, h, x: egnn_ref.egnn_forward(p, y["L"], e, h, x), sch, b.pos, b.x, b.edge_index, b.batch, t,
                                      torch.randn(G * n, 3, generator=g2), torch.randn(G * n, 2, generator=g2), A=2, cond=cond)
            loss_c.backward()
            ro.step()
            if k:
                times.append(time.perf_counter() - t0)
        sec = sum(times) / len(times)
        out["cpu_baseline"] = dict(value=G / sec, unit="graph-steps/s", ms_per_step=1e3 * sec, cores=info["logical_cpus"], kind="port",
                                   cpu_model=info["cpu_model"], sample=f"the same step, {len(times)} timed, oracle/ restatement with torch CPU autograd")
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
