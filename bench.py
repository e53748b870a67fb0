#!/usr/bin/env python
"""Headline benchmark: BASELINE.json configs[1] -- eps-prediction TRAINING step of the EGNN denoiser (6 layers, hidden
128) on a batch of 64 amorphous-SiO2-like 192-atom periodic cutoff graphs (r_c = 5 A) per GPU.

One step = diffuse_as_batch + EGNN forward + loss + backward + optimizer step over one batch (the body of
train_epoch, parts/train_per_iretation.py:122-179).  metric = denoiser graph-steps/s (graphs trained per second,
whole job).  `value` is timed with the batch resident in HBM; `e2e` goes through the public API from pinned HOST
buffers every step (H2D of the batch through data.DevicePrefetcher -- the copy of step i+1 overlaps step i --, edge-index
preparation, D2H of the loss every step, read by the host one step later).  N > 1: one rank per GPU, each its own
64-graph batch (weak scaling), one NCCL all-reduce of the flat gradient per step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

After the training measurement the same line gets a `sampling` object: BASELINE.json configs[2], 512 fully connected
192-atom structures per GPU, T = 1000, through the public batched sampler (structures/s, fraction of the FLOP bound).

--impl reference times the CPU restatement of the reference path (oracle/, torch CPU with all host threads) on the whole
64-graph batch of the same workload; the unmodified reference cannot run here (needs torch_geometric, absent).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import random
import statistics
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(L=6, d=128, H=3, A=2, graphs=64, atoms=192, r_c=5.0, T=1000, s=1e-5, power=2.0, lr=1e-5, seed=2024, name="cfg2",
           cpu_graphs=8)
# --workload cfg4 (BASELINE configs[3]: 1536-atom cells, 9 layers, hidden 256, 8 graphs per GPU); not the contract line.
# Hidden 256 runs the streamed-W2 tcgen05 kernels (tc_edge256.cu, tc_edge256_bwd.cu); --precision fp32 selects the FFMA kernels.
CFG4 = dict(L=9, d=256, graphs=8, atoms=1536, name="cfg4", cpu_graphs=1)
DENSITY = 0.0662  # atoms / A^3, a-SiO2
TRAFFIC_FILE = "r2_traffic.json"  # dram bytes per launch of the top kernels from one `ncu --set full` capture (scripts/profile_round.sh)
# second half of BASELINE.json's metric (configs[2]): x+h ancestral sampling, fully connected 192-atom structures, T = 1000
SAMPLING = dict(structures=512, atoms=192, T=1000, L=6, d=128, H=3, s=0.05, power=2.0)


def synth_batch(num_graphs, n_atoms, seed):
    """Seeded SiO2-like cells: n/3 Si + 2n/3 O uniform in a cubic cell of side (n/0.0662)^(1/3) A with a
    1.4 A periodic min-distance rejection, zero COM.  Returns host tensors pos [N,3], onehot [N,2], box [G,3]."""
    rng = np.random.default_rng(seed)
    side = float((n_atoms / DENSITY) ** (1.0 / 3.0))
    pos_all, oh_all = [], []
    for _ in range(num_graphs):
        pts = np.zeros((n_atoms, 3))
        k = 0
        while k < n_atoms:
            c = rng.uniform(0.0, side, size=3)
            if k:
                d = pts[:k] - c
                d -= side * np.rint(d / side)
                if np.min(np.einsum("ij,ij->i", d, d)) < 1.4 * 1.4:
                    continue
            pts[k] = c
            k += 1
        pos_all.append((pts - pts.mean(axis=0)).astype(np.float32))
        species = np.zeros(n_atoms, np.int64)
        species[rng.permutation(n_atoms)[: n_atoms // 3]] = 1
        oh = np.zeros((n_atoms, 2), np.float32)
        oh[np.arange(n_atoms), species] = 1.0  # [1,0] = O, [0,1] = Si (make_dataset.py:112-117)
        oh_all.append(oh)
    pos = torch.from_numpy(np.concatenate(pos_all))
    onehot = torch.from_numpy(np.concatenate(oh_all))
    box = torch.full((num_graphs, 3), side, dtype=torch.float32)
    return pos, onehot, box


def net_sizes(L, d, H):
    K0 = 2 * H + 1
    return dict(L=L, m_input=K0, m_hidden=d, m_output=d, x_input=K0, x_hidden=d, x_output=1, h_input=H + d, h_hidden=d,
                h_output=H)


def fwd_flops(L, d, H, N, E):
    """Algorithmic forward FLOPs of one call, SURVEY.md 8(d)."""
    K0 = 2 * H + 1
    return L * (E * (4 * K0 * d + 4 * d * d + 4 * d) + N * (2 * (H + d) * d + 2 * d * H))


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled from a thread every 5 ms (short timed
    regions still get samples); `nvidia-smi -lms` as the fallback when the NVML binding is missing."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.nvml, self.stop_flag = [], None, None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        masks = {"hw_slowdown": n.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": n.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": n.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": n.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop_flag:
            try:
                mhz = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                bits = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                self.rows.append((mhz, [k for k, m in masks.items() if bits & m]))
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            if not self.rows:
                return dict(sm_mhz=None, sm_max_mhz=self.max_mhz, reasons=["no samples"], source="nvml")
            return dict(sm_mhz=statistics.median(r[0] for r in self.rows), sm_max_mhz=self.max_mhz,
                        reasons=sorted({k for r in self.rows for k in r[1]}), samples=len(self.rows), source="nvml")
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        rows = [r for r in self.rows if len(r) >= 6 and r[0].isdigit()]
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no samples"])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in rows for k in range(4) if r[2 + k].lower().startswith("active")})
        return dict(sm_mhz=statistics.median(int(r[0]) for r in rows), sm_max_mhz=int(rows[0][1]), reasons=reasons,
                    samples=len(rows), source="nvidia-smi")


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        z = json.load(open(p))
        return dict(hbm_gbs=z["hbm_gbs"], tflops=z.get("bf16_tflops_sustained", z["bf16_tflops"]), source="measured")
    return dict(hbm_gbs=6650.0, tflops=1400.0, source="fallback")


# ------------------------------------------------------------------------------------------- CPU arm (oracle port)
def cpu_train_steps(pos, onehot, edge_index, batch, sd, steps, warmup, threads):
    """Training steps of the CPU restatement (oracle/): egnn_ref forward with autograd, reference loss, torch Adam."""
    from oracle import diffusion_ref as D, egnn_ref, loops_ref as LP
    torch.set_num_threads(threads)
    sch = D.Schedule(CFG["s"], CFG["power"], CFG["T"])
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    opt = torch.optim.Adam(list(params.values()), lr=CFG["lr"])
    G = int(batch.max()) + 1
    gen = torch.Generator().manual_seed(1)
    rnd = random.Random(CFG["seed"])
    times = []
    for k in range(warmup + steps):
        t0 = time.perf_counter()
        t = torch.tensor([rnd.choice(range(1, CFG["T"] + 1)) for _ in range(G)])
        n_pos = torch.randn(pos.shape[0], 3, generator=gen)
        n_h = torch.randn(pos.shape[0], CFG["A"], generator=gen)
        opt.zero_grad()
        loss, _ = LP.train_loss(lambda e, h, x: egnn_ref.egnn_forward(params, CFG["L"], e, h, x), sch, pos, onehot,
                                edge_index, batch, t, n_pos, n_h, A=CFG["A"])
        loss.backward()
        opt.step()
        float(loss.detach())
        if k >= warmup:
            times.append(time.perf_counter() - t0)
    return times


def cpu_sample(pos, onehot, edge_index, node_ptr, graphs):
    """First `graphs` graphs of the batch as CPU tensors (edge list cut at the graph boundary)."""
    n_nodes = int(node_ptr[graphs])
    keep = edge_index[0] < n_nodes
    batch = torch.repeat_interleave(torch.arange(graphs), (node_ptr[1:graphs + 1] - node_ptr[:graphs]).long())
    return pos[:n_nodes].clone(), onehot[:n_nodes].clone(), edge_index[:, keep].clone(), batch


def run_reference(args, rank):
    if rank != 0:
        return
    pos, onehot, box = synth_batch(CFG["graphs"], CFG["atoms"], CFG["seed"] + 1)
    node_ptr = torch.arange(CFG["graphs"] + 1) * CFG["atoms"]
    from oracle import egnn_ref, graphs_ref as G
    torch.manual_seed(CFG["seed"])
    sd = egnn_ref.init_state_dict(**net_sizes(CFG["L"], CFG["d"], CFG["H"]))
    ei = torch.from_numpy(G.cutoff_edges_batch(pos.numpy(), node_ptr.numpy(), CFG["r_c"], box.numpy()))
    cb, total, nsteps = cpu_baseline_block(pos, onehot, ei, node_ptr, sd, args.steps, args.warmup)
    value = cb["value"]
    line = dict(impl="reference", metric="denoiser graph-steps/s (training fwd+bwd+optimizer)", value=value,
                unit="graph-steps/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * total / nsteps, higher_is_better=True, scaling=CFG.get("scaling", "weak"), vs_baseline=None,
                dtype="fp32", data="synthetic", config=workload_config(pos.shape[0], ei.shape[1]),
                cpu_baseline=cb,
                e2e=dict(value=value, unit="graph-steps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                note="CPU restatement of the reference path (oracle/); the unmodified reference needs torch_geometric")
    print(json.dumps(line), flush=True)


def workload_config(N, E):
    """The SAME dict in both arms (the driver compares them); arithmetic type is the top-level `dtype`."""
    return dict(workload=f"{CFG['name']}: eps-pred training step, {CFG['graphs']} x {CFG['atoms']}-atom periodic cutoff graphs (r_c=5A) "
                         f"per GPU, EGNN L={CFG['L']} d={CFG['d']} H=3, T=1000, Adam lr=1e-5", graphs_per_gpu=CFG["graphs"], atoms=CFG["atoms"], r_c=CFG["r_c"], L=CFG["L"],
                hidden=CFG["d"], parallelism="data-parallel, one NCCL all-reduce of the flat gradient per step",
                l2="256 MB buffer written between timed steps (L2 flush, inside the timed region)", nodes_per_gpu=int(N),
                edges_per_gpu=int(E))


def host_info():
    """Host CPU of the box the CPU arm runs on: model name, logical and physical core counts."""
    model = None
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.lower().startswith("model name"):
                    model = line.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    phys = None
    try:
        import psutil
        phys = psutil.cpu_count(logical=False)
    except Exception:
        pass
    return dict(cpu_model=model, logical_cpus=os.cpu_count() or 1, physical_cores=phys or os.cpu_count() or 1)


def cpu_baseline_block(pos, onehot, ei, node_ptr, sd, steps, warmup, budget_s=240.0):
    """The CPU restatement of the reference path (oracle/) timed on the host cores: the WHOLE batch per step with every
    host thread (halved only if one step would blow the time budget), plus a 1-thread and a physical-core figure on an
    8-graph sample.  Returns (cpu_baseline dict, total seconds, timed steps)."""
    info = host_info()
    threads = info["logical_cpus"]
    graphs = CFG["graphs"]
    small = min(CFG["cpu_graphs"], graphs)
    p, o, e, b = cpu_sample(pos, onehot, ei, node_ptr, small)
    probe = cpu_train_steps(p, o, e, b, sd, 1, 1, threads)[0] * graphs / small
    while graphs > 1 and probe * (steps + warmup) > budget_s:
        graphs //= 2
        probe /= 2
    p, o, e, b = cpu_sample(pos, onehot, ei, node_ptr, graphs)
    times = cpu_train_steps(p, o, e, b, sd, steps, warmup, threads)
    value = graphs * len(times) / sum(times)
    ps, os_, es, bs = cpu_sample(pos, onehot, ei, node_ptr, small)
    t1 = cpu_train_steps(ps, os_, es, bs, sd, 1, 0, 1)
    tp = cpu_train_steps(ps, os_, es, bs, sd, 2, 1, info["physical_cores"])
    torch.set_num_threads(threads)
    whole = "all" if graphs == CFG["graphs"] else f"{graphs} of the"
    cb = dict(value=value, unit="graph-steps/s", cores=threads, kind="port",
              sample=f"{whole} {CFG['graphs']} graphs of the batch per step ({p.shape[0]} atoms, {e.shape[1]} edges), {len(times)} timed "
                     f"steps, fp32, oracle/ restatement with torch CPU autograd + Adam",
              cpu_model=info["cpu_model"], logical_cpus=threads, physical_cores=info["physical_cores"],
              value_1_thread=small * len(t1) / sum(t1), value_physical_cores=small * len(tp) / sum(tp),
              thread_sweep_sample=f"{small} graphs per step ({es.shape[1]} edges)")
    return cb, sum(times), len(times)


# ------------------------------------------------------------------------------------------- sampling leg (configs[2])
def tame_(net, L, fx=0.05, fd=0.02):
    """Scale the coordinate head and the distance columns so that an UNTRAINED net gives bounded T-step trajectories (no
    trained weights ship with the reference; default init diverges within a few steps).  Shapes and code path unchanged."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import bump_param_epoch
    with torch.no_grad():
        for l in range(L):
            e = net.egcl_list[l]
            e.mlp_x[4].weight.mul_(fx)
            e.mlp_x[4].bias.mul_(fx)
            e.mlp_m[0].weight[:, -1].mul_(fd)
            e.mlp_x[0].weight[:, -1].mul_(fd)
    bump_param_epoch()


def sampling_leg(args, rank, world, dev, lib, peaks):
    """`structures` x 192-atom fully connected structures per GPU through the public sampler (generate()'s loop batched:
    T ancestral steps + the t = 0 step, COM projection, finite flags), batch-sharded, no collective.  Structures/s is the
    whole job; the bound is the measured bf16 peak over the algorithmic FLOPs of T+1 denoiser calls."""
    import torch.distributed as dist
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.sampler import sample_structures
    c = SAMPLING
    S, n, T, L, d, H = args.sampling_structures, c["atoms"], c["T"], c["L"], c["d"], c["H"]
    torch.manual_seed(CFG["seed"])
    net = EquivariantGNN(**net_sizes(L, d, H)).to(dev)
    tame_(net, L)
    net.precision = args.precision
    sd_cpu = {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    dp = E3DiffusionProcess(s=c["s"], power=c["power"], num_diffusion_timestep=T)
    gen = torch.Generator(device=dev)
    gen.manual_seed(CFG["seed"] + rank)
    sample_structures(net, dp, n, min(S, 4), generator=gen, device=dev)  # warm-up: packs weights, allocates, captures
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    l0 = lib.egnn_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = sample_structures(net, dp, n, S, generator=gen, device=dev)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    ok = out["finite"].sum().to(torch.float32).view(1)
    if world > 1:
        dist.barrier()
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(ok, op=dist.ReduceOp.SUM)
    sec = float(ms.item()) * 1e-3
    flops_structure_step = fwd_flops(L, d, H, n, n * (n - 1))
    bound = peaks["tflops"] * 1e12 / ((T + 1) * flops_structure_step)
    res = dict(metric="sampled structures/s (x+h ancestral sampling, T+1 denoiser calls per structure)",
               value=S * world / sec, unit="structures/s", per_gpu=S / sec, seconds=sec, structures_per_gpu=S,
               accepted_structures=int(ok.item()), denoiser_graph_steps_per_s=S * world * (T + 1) / sec,
               algorithmic_tflops_per_gpu=S * (T + 1) * flops_structure_step / sec / 1e12,
               bound_structures_per_s_per_gpu=bound, frac_of_bound=S / sec / bound, gpu_launches=int(lib.egnn_launch_count() - l0),
               config=dict(workload=f"cfg3: x+h joint diffusion sampling, fully connected {n}-atom structures "
                                    f"(E = {n * (n - 1)} per structure), EGNN L={L} d={d} H=3, T={T}, batch-sharded, no collective",
                           weights="seeded random init, coordinate head x0.05 / distance columns x0.02 (untrained nets diverge)",
                           noise="torch device generator, drawn inside the timed region"))
    if rank == 0 and world == 1:
        # the reference's sampler on the host cores: ONE structure, a bounded number of steps of generate()'s loop
        # (parts/train_per_iretation.py:335-374) through the CPU restatement, extrapolated to T+1 steps
        from oracle import diffusion_ref as D, egnn_ref, graphs_ref as G
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        sch = D.Schedule(c["s"], c["power"], T)
        ei = torch.as_tensor(G.fc_edges(n)).long()
        g2 = torch.Generator().manual_seed(1)
        z_x, z_h = D.remove_mean(torch.randn(n, 3, generator=g2)), torch.randn(n, 2, generator=g2)
        ksteps, times = 12, []
        with torch.no_grad():
            for k in range(ksteps + 2):
                t = T - k
                t0 = time.perf_counter()
                h = torch.cat([z_h, torch.full((n, 1), t / T, dtype=torch.float32)], dim=1)
                nh, nx = egnn_ref.egnn_forward(sd_cpu, L, ei, h, z_x)
                eps_x = D.remove_mean(nx - z_x)
                z_x = D.reverse_diffuse_one_step(sch, z_x, eps_x, t, torch.randn(n, 3, generator=g2), "pos")
                z_h = D.reverse_diffuse_one_step(sch, h[:, :2], nh[:, :2], t, torch.randn(n, 2, generator=g2), "h")
                if k >= 2:
                    times.append(time.perf_counter() - t0)
        per_step = sum(times) / len(times)
        res["cpu_baseline"] = dict(value=1.0 / ((T + 1) * per_step), unit="structures/s", cores=threads, kind="port",
                                   sample=f"1 structure, {ksteps} of the {T + 1} sampler steps ({1e3 * per_step:.1f} ms/step), "
                                          "extrapolated; one structure per call like generate()")
    return res


# ------------------------------------------------------------------------------------------- product arm (B200)
def run_product(args, rank, world, local):
    import torch.distributed as dist
    from diffusion_model_b200 import _lib
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.data import Data, DevicePrefetcher
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.distributed import OverlappedGradSync, broadcast_parameters
    from diffusion_model_b200.graph import build_edges_cutoff, check_pending
    from diffusion_model_b200.optim import FlatAdam
    from diffusion_model_b200.train_per_iretation import GraphedTrainStep, batch_loss

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    lib = _lib.load()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    torch.manual_seed(CFG["seed"])
    random.seed(CFG["seed"] + rank)
    G_, n = CFG["graphs"], CFG["atoms"]
    pos_h, oh_h, box_h = synth_batch(G_, n, CFG["seed"] + 1 + 1000 * rank)
    node_ptr_h = (torch.arange(G_ + 1) * n).to(torch.int32)
    net = EquivariantGNN(**net_sizes(CFG["L"], CFG["d"], CFG["H"])).to(dev)
    net.precision = args.precision
    sd_cpu = {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    broadcast_parameters(net)
    opt = FlatAdam(net, lr=CFG["lr"], capturable=True)
    if world > 1:
        OverlappedGradSync(net, buckets=3)  # gradient all-reduce in 3 buckets of layers, overlapped with backward
    dp = E3DiffusionProcess(s=CFG["s"], power=CFG["power"], num_diffusion_timestep=CFG["T"])
    dp.device_tables(dev)
    prm = dict(atom_type_size=CFG["A"], norm_scope="call")

    # resident batch: edge_index from the CUDA builder (bit-exact against the oracle in tests/)
    pos_d, oh_d, box_d, node_ptr_d = pos_h.to(dev), oh_h.to(dev), box_h.to(dev), node_ptr_h.to(dev)
    ei_d = build_edges_cutoff(pos_d, CFG["r_c"], node_ptr=node_ptr_d, box=box_d)
    N, E = pos_d.shape[0], ei_d.shape[1]
    resident = Data(x=oh_d, pos=pos_d, edge_index=ei_d)
    resident.node_ptr = node_ptr_d
    ei_h = ei_d.cpu().pin_memory()
    pos_p, oh_p, np_p = pos_h.pin_memory(), oh_h.pin_memory(), node_ptr_h.pin_memory()

    def train_step(b):
        """the step enqueued kernel by kernel through Python (what train_epoch does for minibatches of varying shape)"""
        opt.zero_grad()
        loss, _, _ = batch_loss(net, b, dp, prm)
        loss.backward()   # N > 1: the bucketed all-reduce runs on the sync's stream under the rest of backward
        opt.step(grad_scale=1.0 / world)   # waits for the last bucket
        return loss

    # e2e: the public loader path.  Every step's inputs start in pinned HOST memory; DevicePrefetcher issues step
    # i+1's H2D copy on a side stream while step i computes (one copy per step inside the timed region), and the
    # step's loss is copied back to pinned memory every step and read by the host one step later (no stall).
    host_batch = Data(x=oh_p, pos=pos_p, edge_index=ei_h)
    host_batch.node_ptr = np_p

    def host_batches():
        while True:
            yield host_batch

    # The product's fast path for fixed-shape minibatches: the whole step captured once in a CUDA graph
    # (train_per_iretation.GraphedTrainStep); `--eager` times the kernel-by-kernel path instead.
    graphed = None if args.eager else GraphedTrainStep(net, dp, prm, opt, resident, grad_scale=1.0 / world, device=dev)
    feed = iter(DevicePrefetcher(host_batches(), dev))
    loss_pin = torch.zeros(2, pin_memory=True)
    loss_ev = [torch.cuda.Event(), torch.cuda.Event()]
    e2e_state = dict(i=0, last=None)

    def e2e_step():
        if graphed is None:
            b = next(feed)
            loss = train_step(b)
        else:
            if e2e_state["i"] == 0:
                graphed.prefetch(host_batch)
            loss = graphed.step()            # moves the staged minibatch into the static buffers, replays
            graphed.prefetch(host_batch)     # H2D of the NEXT step's inputs on the copy stream, under this replay
        k = e2e_state["i"] & 1
        loss_pin[k:k + 1].copy_(loss.detach().view(1), non_blocking=True)  # D2H read of the step's loss
        loss_ev[k].record()
        if e2e_state["i"] > 0:
            loss_ev[k ^ 1].synchronize()
            e2e_state["last"] = float(loss_pin[k ^ 1])
            if not math.isfinite(e2e_state["last"]):
                raise SystemExit("bench.py: non-finite loss in the e2e loop")
        e2e_state["i"] += 1
        check_pending(wait=False)   # deferred edge-list validation of the graphs whose flags have landed

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        clocks = ClockSampler(local)
        l0 = lib.egnn_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            flush.zero_()
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), lib.egnn_launch_count() - l0, clocks.stop()

    # kernels one step launches (the library counts its launches on the host; a graph replay does not pass there)
    torch.cuda.synchronize(dev)
    l0 = lib.egnn_launch_count()
    train_step(resident)
    torch.cuda.synchronize(dev)
    launches_per_step = int(lib.egnn_launch_count() - l0)
    resident_step = (lambda: train_step(resident)) if graphed is None else (lambda: graphed.step())
    ms, _, clocks = timed(resident_step, args.steps, args.warmup)
    launches = launches_per_step * args.steps
    value = G_ * world * args.steps / (ms * 1e-3)
    eager = None
    if graphed is not None:
        k_eager = min(args.steps, 10)
        ms_eager, _, _ = timed(lambda: train_step(resident), k_eager, 3)
        eager = dict(ms_per_step=ms_eager / k_eager, value=G_ * world * k_eager / (ms_eager * 1e-3),
                     note="same step enqueued kernel by kernel through Python (no CUDA graph)")
    ms_e2e, _, _ = timed(e2e_step, args.steps, max(args.warmup, 3))
    e2e_value = G_ * world * args.steps / (ms_e2e * 1e-3)
    h2d = pos_p.numel() * 4 + oh_p.numel() * 4 + ei_h.numel() * 8 + G_ * 4 + (np_p.numel() * 4 if graphed is None else 0)
    # host time to enqueue one step (no sync inside): the floor of the step time when the GPU is not the bottleneck
    host_ms = []
    for _ in range(5):
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        resident_step()
        host_ms.append(1e3 * (time.perf_counter() - t0))
    torch.cuda.synchronize(dev)
    host_ms = statistics.median(host_ms)

    # per-kernel device times (CUDA events on the launching stream, separate short pass after the timed regions)
    import ctypes as C
    lib.egnn_profile_enable(1)
    prof_steps = 3
    for _ in range(prof_steps):
        flush.zero_()
        train_step(resident)
    torch.cuda.synchronize(dev)
    names = {0: "egcl_edge_fwd", 1: "egcl_node_fwd", 2: "egcl_xupdate", 3: "egcl_edge_bwd", 4: "egcl_node_bwd",
             5: "egcl_gather_bwd", 7: "tc_edge_fwd", 8: "tc_edge_bwd", 9: "tc_wgrad"}
    prof = {}
    for kind, name in names.items():
        tot, cnt = C.c_double(0), C.c_int64(0)
        lib.egnn_profile_read(kind, C.byref(tot), C.byref(cnt))
        if cnt.value:
            prof[name] = dict(ms_per_launch=tot.value / cnt.value, launches_per_step=cnt.value / prof_steps,
                              ms_per_step=tot.value / prof_steps)
    lib.egnn_profile_enable(0)
    K0, d = 2 * CFG["H"] + 1, CFG["d"]
    edge_flops = E * (4 * K0 * d + 4 * d * d + 4 * d)  # algorithmic forward FLOPs of the edge MLPs of one layer
    algo = {"egcl_edge_fwd": edge_flops, "tc_edge_fwd": edge_flops, "egcl_edge_bwd": 2 * edge_flops,
            "tc_edge_bwd": 2 * edge_flops}
    peaks = measured_peaks()
    top = max(prof, key=lambda k: prof[k]["ms_per_step"]) if prof else None
    roofline = None
    if top in algo:
        # d = 256: the edge backward is TWO kernels per layer (tc_edge_bwd256 writes the tile images, tc_wgrad256 turns
        # them into dW2): the backward FLOPs are divided by the time of the pair
        pair_ms = prof[top]["ms_per_launch"]
        if top == "tc_edge_bwd" and "tc_wgrad" in prof:
            pair_ms += prof["tc_wgrad"]["ms_per_launch"]
        ach = algo[top] / (pair_ms * 1e-3) / 1e12
        # tensor-core FLOPs actually issued: the backward recomputes both edge MLPs (3x forward), and bf16x3 issues
        # three bf16 MMAs (hi.hi + lo.hi + hi.lo) per fp32-accurate product
        # (forward: GEMM1 + GEMM2 of both branches; backward: the same recomputed + dgrad, and the weight-gradient MMAs
        # with two of the three split products)
        passes = {"fp32": 0, "bf16x3": 3, "bf16": 1}[args.precision] if top.startswith("tc_") else 0
        wgrad_passes = min(passes, 2)
        issued = (2 * passes + wgrad_passes if top.endswith("bwd") else passes) * edge_flops
        traffic = None
        try:
            tkey = top + ("256" if CFG["d"] == 256 else "")  # ncu captures: cfg2 for the d=128 kernels, cfg4 for d=256
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", TRAFFIC_FILE)) as f:
                traffic = json.load(f).get(tkey, {}).get("dram_bytes_per_launch")
        except OSError:
            try:
                with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as f:
                    traffic = json.load(f).get(tkey, {}).get("dram_bytes_per_launch")
            except OSError:
                pass
        roofline = dict(bound="tensor", kernel=top, achieved=ach, peak=peaks["tflops"], unit="TFLOP/s",
                        frac=ach / peaks["tflops"], traffic=traffic, peak_source=peaks["source"] + " bf16 sustained",
                        share_of_step=pair_ms * prof[top]["launches_per_step"] / (ms / args.steps),
                        flops_per_launch=algo[top], ms_per_launch=pair_ms,
                        issued_tensor_flops_per_launch=issued,
                        issued_frac=issued / (pair_ms * 1e-3) / 1e12 / peaks["tflops"])
    line = dict(metric="denoiser graph-steps/s (training fwd+bwd+optimizer)", value=value, unit="graph-steps/s",
                n_gpus=world, steps=args.steps, warmup=args.warmup, ms_per_step=ms / args.steps, higher_is_better=True,
                scaling=CFG.get("scaling", "weak"), vs_baseline=None, dtype=args.precision, data="synthetic",
                config=workload_config(N, E),
                e2e=dict(value=e2e_value, unit="graph-steps/s", h2d_bytes_per_step=h2d, d2h_bytes_per_step=4,
                         pipeline="H2D of step i+1 staged on a copy stream during step i (then one device-to-device move into the "
                                  "graph's static buffers); loss D2H every step, read 1 step late",
                         ms_per_step=ms_e2e / args.steps, host_enqueue_ms_per_step=host_ms),
                gpu_launches=int(launches), launches_per_step=launches_per_step,
                step_mode=("eager: kernels enqueued one by one" if graphed is None else
                           "CUDA graph: the whole step (diffusion, forward, loss, backward, all-reduce, Adam) captured once and "
                           "replayed, a new minibatch's edge-list preparation in front of it; gpu_launches = kernels per step x steps"),
                eager=eager, clocks=clocks, roofline=roofline, kernels=prof,
                train_flops_per_step=3 * fwd_flops(CFG["L"], d, CFG["H"], N, E))
    if rank == 0 and world == 1:
        cb, _, _ = cpu_baseline_block(pos_h, oh_h, ei_d.cpu(), node_ptr_h.long(), sd_cpu, 6, 1, budget_s=30.0)
        line["cpu_baseline"] = cb
    if args.sampling_structures > 0 and CFG["name"] == "cfg2":
        line["sampling"] = sampling_leg(args, rank, world, dev, lib, peaks)
    if rank == 0:
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default="bf16x3", choices=["fp32", "bf16x3", "bf16"],
                    help="bf16x3 (default): tcgen05 split-bf16 products, fp32 accumulate, <=1e-4 per layer vs the fp32 "
                         "reference; fp32: CUDA-core FFMA; bf16: single tcgen05 product (looser, stated bound)")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="strong scaling: this many graphs in total, split evenly over the ranks (default 0: weak scaling, "
                         "the workload's graph count on EVERY rank)")
    ap.add_argument("--eager", action="store_true", help="time the kernel-by-kernel step instead of the CUDA-graph replay")
    ap.add_argument("--sampling-structures", type=int, default=SAMPLING["structures"],
                    help="structures per GPU of the sampling leg (cfg3: 512 x 192 atoms fully connected, T = 1000); 0 skips it")
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg4", "yaml"],
                    help="cfg2 (default, the contract line), cfg4 (1536-atom cells, L=9, hidden 256) or yaml (the reference's own "
                         "parameters.yaml network, one GPU: bench_yaml.py; --global-batch = graphs per step, default 1)")
    args = ap.parse_args()
    if args.workload == "yaml":
        import bench_yaml
        if args.impl != "b200":
            raise SystemExit("--workload yaml has no separate reference arm: its CPU oracle timing is the cpu_baseline block of the "
                             "product line (bench_yaml.py); the reference arm of the contract runs the default workload")
        bench_yaml.main(["--steps", str(args.steps), "--batch", str(args.global_batch or 1)])
        return
    if args.workload == "cfg4":
        CFG.update(CFG4)
    CFG["scaling"] = "weak"
    if args.global_batch:
        world_ = int(os.environ.get("WORLD_SIZE", "1"))
        if args.global_batch % world_:
            raise SystemExit(f"--global-batch {args.global_batch} is not a multiple of the {world_} ranks")
        CFG["graphs"] = args.global_batch // world_
        CFG["cpu_graphs"] = min(CFG["cpu_graphs"], CFG["graphs"])
        CFG["scaling"] = "strong"
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args, int(os.environ.get("RANK", "0")))
        return
    from diffusion_model_b200.distributed import init_from_env
    rank, world, local = init_from_env("nccl")
    try:
        run_product(args, rank, world, local)
    finally:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
