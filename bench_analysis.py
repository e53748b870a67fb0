"""Structure-analysis throughput: the "evaluate" half of sample -> evaluate at the cfg3 size,
4096 sampled structures x 192 atoms.  Times egnn_rdf_local (histogram + normalisation + Gaussian smoothing),
egnn_cn2_sites and egnn_kabsch with CUDA events and reports each against the HBM roofline (algorithmic bytes: the
position read, 12 B per atom, plus the per-structure outputs), next to the reference's python loops restated in
oracle/eval_ref.py timed on a bounded sample of the same structures (cpu_baseline).  Prints ONE JSON line.

    python bench_analysis.py [--structures 4096] [--atoms 192] [--iters 20]
"""
from __future__ import annotations

import argparse
import json
import os
import time

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--structures", type=int, default=4096)
    ap.add_argument("--atoms", type=int, default=192)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--cpu-sample", type=int, default=16)
    args = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("bench_analysis.py: no CUDA device; the product has no CPU path")
    from diffusion_model_b200 import _lib, evaluate as EV
    lib = _lib.load()
    S, n = args.structures, args.atoms
    g = torch.Generator().manual_seed(2024)
    host = 2.5 * torch.randn(S, n, 3, generator=g)
    pos = host.cuda()
    ref = (host + 0.05 * torch.randn(S, n, 3, generator=g)).cuda()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    peaks = {}
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    hbm = float(peaks.get("hbm_gbs", 6500.0))

    def timed(fn):
        for _ in range(3):
            fn()
        tot = 0.0
        for _ in range(args.iters):
            flush.fill_(1)  # 256 MB > 126 MB L2: the positions come from HBM
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            tot += a.elapsed_time(b)
        return tot / args.iters

    nb = 500
    l0 = lib.egnn_launch_count()
    res = {}
    for name, fn, bytes_ in (
            ("rdf_local", lambda: EV.RDF_batch(pos, return_counts=True), S * n * 12 + S * nb * 12),
            ("cn2_sites", lambda: EV.cn2_sites(pos, radius=2.0), S * n * 12 + S * 28),
            ("kabsch", lambda: EV.kabsch_batch(pos, ref), 3 * 2 * S * n * 12 + S * 52)):  # P and Q are read three times
        ms = timed(fn)
        res[name] = dict(ms=ms, structures_per_s=S / ms * 1e3, bytes=bytes_, gbs=bytes_ / ms / 1e6,
                         hbm_frac=bytes_ / ms / 1e6 / hbm)
    launches = lib.egnn_launch_count() - l0
    # reference python loops (oracle restatement) on a bounded sample
    from oracle import eval_ref as E
    k = min(args.cpu_sample, S)
    t0 = time.perf_counter()
    for s in range(k):
        E.rdf(host[s])
    t_rdf = (time.perf_counter() - t0) / k
    t0 = time.perf_counter()
    for s in range(k):
        E.cn2_site(host[s], 2.0)
        E.kabsch(host[s], ref[s].cpu())
    t_rest = (time.perf_counter() - t0) / k
    total_ms = sum(r["ms"] for r in res.values())
    print(json.dumps({
        "metric": "evaluated structures/s (local RDF + CN2 site + Kabsch RMSD)", "value": S / total_ms * 1e3,
        "unit": "structures/s", "n_gpus": 1, "dtype": "f32 distances, i32 counts, f64 curves", "data": "synthetic",
        "config": {"workload": f"{S} structures x {n} atoms, R=5 dR=0.01 sigma=5 (500 bins), radius 2.0", "l2": "256 MB flush before every timed launch"},
        "kernels": res, "gpu_launches": launches,
        "roofline": {"bound": "hbm", "kernel": "rdf_local_kernel", "achieved": res["rdf_local"]["gbs"], "peak": hbm,
                     "unit": "GB/s", "frac": res["rdf_local"]["hbm_frac"], "traffic": None},
        "cpu_baseline": {"value": 1.0 / (t_rdf + t_rest), "unit": "structures/s", "cores": 1, "kind": "port",
                         "sample": f"{k} of the {S} structures, oracle/eval_ref.py (the reference's python loops)",
                         "rdf_ms_per_structure": t_rdf * 1e3, "cn2_kabsch_ms_per_structure": t_rest * 1e3}}))


if __name__ == "__main__":
    main()
