"""TEST INFRASTRUCTURE ONLY -- freezes structure sets drawn by the CPU ORACLE sampler (oracle/loops_ref.py, the restated
generate() loop of parts/train_per_iretation.py:299-428 over the restated EGNN) as tests/golden/sampling_<set>.npz, so that
the distribution-level GPU test (tests/test_gpu_sampling_dist.py) compares the CUDA sampler against a LARGE oracle sample
without paying minutes of CPU sampling per test run.  Deterministic: seeded weights, seeded noise.

    python -m oracle.make_sampling_golden [set24] [set192]      (about 15 minutes on 8 cores for both)

  set24   : 512 structures x 24 atoms, EGNN L=4 d=64, T=250, s=0.05   (two noise seeds x 256)
  set192  : 24 structures x 192 atoms fully connected (cfg3's shape), EGNN L=6 d=128, T=40, s=0.2, stronger taming
            (fx=0.005, fd=0.002: with 191 neighbours per atom and 25x longer steps than T=1000 the default taming diverges --
            at fx=0.01, fd=0.005 the oracle's 24 trajectories stayed finite but a few of 192 on the GPU did not;
            s=0.2 keeps the growth of the structure scale, 1/alpha_T, at 5x so that the d^2 feature stays bounded)
Each set stores the sum / abs-sum of its weights so that the test can assert it rebuilt the same network.
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import diffusion_ref as D, egnn_ref, loops_ref as LP  # noqa: E402

SETS = dict(
    set24=dict(L=4, d=64, n=24, T=250, s=0.05, wseed=0, fx=0.05, fd=0.02, noise_seeds=(11, 12), per_seed=256),
    set192=dict(L=6, d=128, n=192, T=40, s=0.2, wseed=1, fx=0.005, fd=0.002, noise_seeds=(21, 22), per_seed=12),
)


def weights_of(c):
    torch.manual_seed(c["wseed"])
    sizes = egnn_ref.net_sizes(c["L"], c["d"])
    return sizes, egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), c["L"], fx=c["fx"], fd=c["fd"])


def weight_checksum(sd):
    return np.array([sum(float(v.double().sum()) for v in sd.values()), sum(float(v.double().abs().sum()) for v in sd.values())])


def main():
    for name in (sys.argv[1:] or list(SETS)):
        c, out = SETS[name], {}
        _, sd = weights_of(c)
        sch = D.Schedule(c["s"], 2.0, c["T"])
        chunks = []
        for seed in c["noise_seeds"]:
            t0 = time.time()
            gen = torch.Generator().manual_seed(seed)
            per = [LP.make_noises(c["n"], c["T"], 2, gen) for _ in range(c["per_seed"])]
            cat = {k: torch.cat([p[k] for p in per], dim=1 if k.startswith("step") else 0) for k in per[0]}
            with torch.no_grad():
                pos, _, fin = LP.sample_structures_batched(sd, c["L"], sch, c["n"], c["per_seed"], cat)
            assert bool(fin.all()), (name, seed)
            chunks.append(pos)
            print(name, "seed", seed, "%.0f s" % (time.time() - t0), flush=True)
        out["pos"] = torch.cat(chunks).numpy().astype(np.float32)
        out["wsum"] = weight_checksum(sd)
        path = os.path.join(ROOT, "tests", "golden", f"sampling_{name}.npz")
        np.savez_compressed(path, **out)
        print("wrote", path, {k: v.shape for k, v in out.items()}, flush=True)


if __name__ == "__main__":
    main()
