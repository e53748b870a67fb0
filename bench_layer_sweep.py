"""EGCL layer microbenchmark sweep (BASELINE configs[4]): one EGCL layer on random receiver-sorted graphs,
atoms 64-4096, average degree 16-128, hidden 64-256, fp32 (CUDA-core FFMA) vs the tcgen05 modes, forward and
forward+backward, with the fraction of the measured B200 bf16 tensor peak and of the HBM peak next to each line.

    python bench_layer_sweep.py [--quick] > profiles/r1_layer_sweep.jsonl

One JSON object per line.  FLOPs / bytes are the algorithmic figures of SURVEY.md 8(d):
    F_layer = E (4 K0 d + 4 d^2 + 4 d) + N (2 d^2 + 4 H d),   B_layer = 8 N (H + 3) + 8 E + 4 P_layer.
Hidden 256 runs the streamed-W2 tcgen05 kernels (tc_edge256.cu forward, tc_edge256_bwd.cu backward + wgrad GEMM); its node
MLP stays on the FFMA kernels.
"""
from __future__ import annotations

import argparse
import json
import os

import torch


def random_graph(n, deg, gen, dev):
    """`deg` distinct random senders per node (no self loops), receiver-sorted, senders ascending per receiver"""
    deg = min(deg, n - 1)
    scores = torch.rand(n, n, generator=gen)
    scores.fill_diagonal_(2.0)
    send = torch.sort(torch.topk(scores, deg, dim=1, largest=False).indices, dim=1).values
    recv = torch.arange(n).view(n, 1).expand(n, deg)
    return torch.stack([recv.reshape(-1), send.reshape(-1)]).to(dev)


def peaks():
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")
    try:
        z = json.load(open(path))
        return z.get("bf16_tflops_sustained", z["bf16_tflops"]), z["hbm_gbs"], "measured"
    except (OSError, KeyError):
        return 1400.0, 6650.0, "fallback"


def time_ms(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(iters):
        flush.zero_()  # L2 flush outside the timed interval
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--hidden", type=int, nargs="*", default=None)
    args = ap.parse_args()
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EGCL
    from diffusion_model_b200.graph import GraphIndex
    dev = torch.device("cuda")
    tf_peak, hbm_peak, src = peaks()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    gen = torch.Generator().manual_seed(2024)
    H = 3
    K0 = 2 * H + 1
    atoms = [256, 4096] if args.quick else [64, 256, 1024, 4096]
    degs = [32] if args.quick else [16, 32, 64, 128]
    hidden = args.hidden or ([128] if args.quick else [64, 128, 256])
    for d in hidden:
        torch.manual_seed(0)
        layer = EGCL(K0, d, d, K0, d, 1, H + d, d, H).to(dev)
        P = sum(p.numel() for p in layer.parameters())
        modes = ["fp32", "bf16x3", "bf16"]  # d = 256: streamed-W2 kernels (tc_edge256.cu, tc_edge256_bwd.cu)
        for n in atoms:
            for deg in degs:
                if deg >= n:
                    continue
                ei = random_graph(n, deg, gen, dev)
                E = ei.shape[1]
                gi = GraphIndex(ei, n)
                h = torch.randn(n, H, generator=gen).to(dev)
                x = torch.randn(n, 3, generator=gen).to(dev)
                flops = E * (4 * K0 * d + 4 * d * d + 4 * d) + n * (2 * d * d + 4 * H * d)
                nbytes = 8 * n * (H + 3) + 8 * E + 4 * P
                for mode in modes:
                    layer.precision = mode

                    def fwd():
                        with torch.no_grad():
                            layer(ei, h, x, graph=gi)

                    hg, xg = h.clone().requires_grad_(True), x.clone().requires_grad_(True)

                    def fwd_bwd():
                        layer.zero_grad(set_to_none=True)
                        ho, xo = layer(ei, hg, xg, graph=gi)
                        (ho.sum() + xo.sum()).backward()

                    ms_f = time_ms(fwd, args.iters, flush)
                    ms_fb = time_ms(fwd_bwd, args.iters, flush)
                    print(json.dumps(dict(
                        atoms=n, degree=deg, edges=E, hidden=d, precision=mode,
                        fwd_us=1e3 * ms_f, fwd_bwd_us=1e3 * ms_fb,
                        fwd_tflops=flops / (ms_f * 1e-3) / 1e12, fwd_bwd_tflops=3 * flops / (ms_fb * 1e-3) / 1e12,
                        fwd_frac_tensor_peak=flops / (ms_f * 1e-3) / 1e12 / tf_peak,
                        fwd_bwd_frac_tensor_peak=3 * flops / (ms_fb * 1e-3) / 1e12 / tf_peak,
                        fwd_gbs=nbytes / (ms_f * 1e-3) / 1e9, fwd_frac_hbm_peak=nbytes / (ms_f * 1e-3) / 1e9 / hbm_peak,
                        flops_fwd=flops, bytes_fwd=nbytes, peaks=dict(tflops=tf_peak, hbm_gbs=hbm_peak, source=src),
                        note=None if d in (64, 128) or mode == "fp32" else "hidden 256: W2 streamed through shared memory, node MLP on the FFMA kernels")), flush=True)


if __name__ == "__main__":
    main()
