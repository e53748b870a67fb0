"""Distribution-level parity of the sampler (SURVEY.md 8c "Tolerances", BASELINE north_star): structures drawn by the
CUDA sampler and by the CPU oracle with the SAME weights and schedule but independent noise must be indistinguishable
under the reference's structure metrics -- the local RDF around atom 0 (evaluate_RDF.py:39-63, restated and pinned in
oracle/eval_ref.py) and the angle / bond lengths of atom 0 with two neighbours (CN2_evaluate.py:12-21).

The oracle side is a frozen fixture: oracle/make_sampling_golden.py sampled 512 x 24-atom structures (EGNN L=4 d=64, T=250)
and 24 x 192-atom structures (cfg3's shape, L=6 d=128, T=40) on the CPU and stored them under tests/golden/ -- minutes of
CPU sampling that buy a large oracle sample without costing every test run.  The test rebuilds the same seeded network
(checked against the fixture's weight checksum), samples on the GPU, and compares.

What is compared, and why (measured on two independent ORACLE sets, i.e. the metrics' own noise floor):
  * mean local-RDF curve, cosine similarity >= 0.99 (SURVEY's bar).  R is the 99th percentile of the oracle's distances
    from atom 0 and the curve has 60 bins: with R taken from the largest distance of the pooled sets (heavy tail) and 480
    bins the SAME sampler scores only 0.93-0.99 against itself at 128-256 structures per side; at 256 per side with
    these settings it scores 0.9996.
  * two-sample Kolmogorov-Smirnov on PER-STRUCTURE statistics (mean and smallest distance from atom 0, radius of gyration,
    CN2 angle and mean bond length of atom 0 with its two nearest atoms): one value per structure, so the samples are
    independent.  (The 23 distances of one structure are correlated; pooled, they make a KS test reject the oracle
    against itself.)  Bar p > 0.01; seeds are fixed, so the outcome is deterministic.
No trained weights ship with the reference, so the nets are tamed random inits; their structures have no 2-Angstrom shell,
hence the two NEAREST atoms stand in for the radius rule of evaluate_Si-O-Si.py:28-33.
"""
import os

import numpy as np
import pytest
import torch
from scipy.stats import ks_2samp

from oracle import eval_ref as E
from oracle import make_sampling_golden as MG

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _fixture(golden_dir, name):
    z = np.load(os.path.join(golden_dir, f"sampling_{name}.npz"))
    c = MG.SETS[name]
    sizes, sd = MG.weights_of(c)
    assert np.allclose(MG.weight_checksum(sd), z["wsum"], rtol=1e-12, atol=0), "rebuilt weights differ from the fixture's"
    return torch.from_numpy(z["pos"]), c, sizes, sd


def _gpu_set(c, sizes, sd, S, precision, seed, min_accept=1.0):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.sampler import sample_structures
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    net.precision = precision
    dp = E3DiffusionProcess(s=c["s"], power=2.0, num_diffusion_timestep=c["T"])
    g = torch.Generator(device=DEV)
    g.manual_seed(seed)
    out = sample_structures(net, dp, c["n"], S, generator=g)
    # generate() rejects non-finite structures and redraws (parts/train_per_iretation.py:376-389): keep the accepted ones;
    # an untrained net may lose the odd trajectory, a broken sampler would lose many
    ok = out["finite"].cpu()
    assert float(ok.float().mean()) >= min_accept, f"only {int(ok.sum())} of {S} structures finite"
    return out["pos"].cpu()[ok]


def _d0(s):
    return (s[:, 1:] - s[:, :1]).norm(dim=2)  # [S, n-1] distances from atom 0


def _per_structure(s):
    d = _d0(s)
    rg = (s - s.mean(dim=1, keepdim=True)).norm(dim=2).pow(2).mean(dim=1).sqrt()
    angles, bonds = [], []
    for p in s:
        tri = p[E.two_nearest(p)]
        angles.append(E.angle_cn2(tri))
        b1, b2 = E.bond_lengths_cn2(tri)
        bonds.append(0.5 * (b1 + b2))
    return dict(mean_dist=d.mean(dim=1).numpy(), min_dist=d.min(dim=1).values.numpy(), radius_of_gyration=rg.numpy(),
                cn2_angle=np.array(angles), cn2_bond=np.array(bonds))


def _compare(pos_c, pos_g, nbins, tag):
    R = float(torch.quantile(_d0(pos_c).flatten(), 0.99))
    dR = R / nbins
    sc = E.structure_set_stats(list(pos_c), R, dR)
    sg = E.structure_set_stats(list(pos_g), R, dR)
    cos = E.cos_similarity(sc["rdf"], sg["rdf"])
    wd = E.wasserstein(sc["rdf"], sg["rdf"])
    pc, pg = _per_structure(pos_c), _per_structure(pos_g)
    ps = {k: float(ks_2samp(pc[k], pg[k]).pvalue) for k in pc}
    print(f"{tag}: {len(pos_c)} oracle vs {len(pos_g)} CUDA structures | RDF cos-sim {cos:.4f} wasserstein {wd:.3e} | KS p "
          + " ".join(f"{k} {v:.3f}" for k, v in ps.items()))
    return cos, ps


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_sampled_structure_distributions_match_oracle(lib, golden_dir, precision):
    pos_c, c, sizes, sd = _fixture(golden_dir, "set24")
    pos_g = _gpu_set(c, sizes, sd, 2048, precision, seed=12)
    cos, ps = _compare(pos_c, pos_g, 60, f"24 atoms T={c['T']} {precision}")
    assert cos >= 0.99, cos
    assert min(ps.values()) > 0.01, ps


def test_sampled_distributions_at_cfg3_shape(lib, golden_dir):
    """The same check on cfg3's shape -- 192-atom fully connected structures, EGNN L=6 d=128 (the sampling bench's net) --
    with a short schedule (T = 40).  24 oracle structures: 4 584 distances for the RDF curve (30 bins), and per-structure
    statistics of 24 against 192."""
    pos_c, c, sizes, sd = _fixture(golden_dir, "set192")
    pos_g = _gpu_set(c, sizes, sd, 192, "bf16x3", seed=22, min_accept=0.97)
    cos, ps = _compare(pos_c, pos_g, 30, f"192 atoms T={c['T']} bf16x3")
    assert cos >= 0.99, cos
    assert min(ps.values()) > 0.01, ps


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_cfg3_shape_layers_match_oracle(lib, precision):
    """Deterministic parity at cfg3's shape: 8 fully connected 192-atom structures in one call (E = 293 376; every receiver
    owns 191 edges, so every receiver run crosses 128-edge tile boundaries), EGNN L=6 d=128, normaliser per structure.
    Each layer is run on the ORACLE's own layer input and compared with the oracle's output: 1e-4 per layer.

    Why per layer and not per trajectory: with untrained weights this net amplifies round-off by ~300x per denoiser call
    (no residual or normalisation on h, sums over 191 messages).  Measured: the fp32 ORACLE against the float64 oracle is
    0.29 apart after THREE calls on one of eight structures (4e-3 .. 0.26 after 5 .. 17 calls); on the GPU the fp32 and
    bf16x3 runs of one T = 40 trajectory end 0.1-1.0 apart on all 24 structures of the fixture.  Equality of multi-call
    trajectories under identical noise is not a well-posed test at this shape; the distribution test above covers the
    long horizon, this one pins the arithmetic of every call."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.graph import build_edges_fc
    from oracle import egnn_ref
    c = dict(MG.SETS["set192"])
    S, n, L = 8, c["n"], c["L"]
    sizes, sd = MG.weights_of(c)
    gen = torch.Generator().manual_seed(41)
    x0 = torch.randn(S * n, 3, generator=gen)
    x0 = (x0.view(S, n, 3) - x0.view(S, n, 3).mean(dim=1, keepdim=True)).reshape(S * n, 3)
    h0 = torch.cat([torch.randn(S * n, 2, generator=gen), torch.full((S * n, 1), 0.5)], dim=1)
    ei = build_edges_fc([n] * S, DEV)
    batch = torch.arange(S).repeat_interleave(n)
    with torch.no_grad():
        _, _, layers = egnn_ref.egnn_forward(sd, L, ei.cpu(), h0, x0, "graph", batch=batch, return_layers=True)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    node_ptr = (torch.arange(S + 1) * n).to(torch.int32).to(DEV)
    rel = lambda a, b: float((a.detach().cpu().double() - b.double()).abs().max() / b.double().abs().max())
    h, x, worst = h0, x0, (0.0, 0.0)
    with torch.no_grad():
        for l, (hr, xr) in enumerate(layers):
            layer = net.egcl_list[l]
            layer.precision = precision
            ho, xo = layer(ei, h.to(DEV), x.to(DEV), node_ptr=node_ptr, norm_scope="graph")
            eh, ex = rel(ho, hr), rel(xo, xr)
            assert eh <= 1e-4 and ex <= 1e-4, (l, eh, ex)
            worst = (max(worst[0], eh), max(worst[1], ex))
            h, x = hr, xr
    print(f"cfg3 shape per layer, {precision}: worst h {worst[0]:.2e} x {worst[1]:.2e} (|h| grows to {float(layers[-1][0].abs().max()):.1e})")
