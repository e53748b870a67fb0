"""Distribution-level parity of the sampler (SURVEY.md 8c "Tolerances", BASELINE north_star): structures drawn by the
CUDA sampler and by the CPU oracle with the SAME weights and schedule but independent noise must be indistinguishable
under the reference's structure metrics -- the local RDF around atom 0 (evaluate_RDF.py:39-63, restated and pinned in
oracle/eval_ref.py) and the angle / bond lengths of atom 0 with two neighbours (CN2_evaluate.py:12-21).

No trained weights ship with the reference, so the net is the tamed random init used by the trajectory fixture; its
structures have no 2-Angstrom shell, hence R/dR are scaled to the sample's extent and the two NEAREST atoms stand in for
the radius rule of evaluate_Si-O-Si.py:28-33.  Bars: cosine similarity of the mean RDF curves >= 0.99, two-sample
Kolmogorov-Smirnov p > 0.01 for distances, angles and bond lengths (seeds are fixed, so the outcome is deterministic).
"""
import numpy as np
import pytest
import torch
from scipy.stats import ks_2samp

from oracle import diffusion_ref as D
from oracle import egnn_ref, eval_ref as E, loops_ref as LP

pytestmark = pytest.mark.gpu
DEV = "cuda"


_ORACLE = {}


def _oracle_set(sd, L, sch, n, T, S):
    if "pos" not in _ORACLE:  # one CPU run serves both precisions
        gen = torch.Generator().manual_seed(11)
        per = [LP.make_noises(n, T, 2, gen) for _ in range(S)]
        cat = {k: torch.cat([p[k] for p in per], dim=1 if k.startswith("step") else 0) for k in per[0]}
        with torch.no_grad():
            _ORACLE["pos"], _, _ORACLE["fin"] = LP.sample_structures_batched(sd, L, sch, n, S, cat)
    return _ORACLE["pos"], _ORACLE["fin"]


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_sampled_structure_distributions_match_oracle(lib, precision):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.sampler import sample_structures
    torch.manual_seed(0)
    L, d, n, T, S_cpu, S_gpu = 4, 64, 24, 250, 256, 2048
    sizes = egnn_ref.net_sizes(L, d)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), L)
    sch = D.Schedule(0.05, 2.0, T)
    pos_c, fin_c = _oracle_set(sd, L, sch, n, T, S_cpu)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    net.precision = precision
    dp = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=T)
    g = torch.Generator(device=DEV)
    g.manual_seed(12)
    out = sample_structures(net, dp, n, S_gpu, generator=g)
    assert bool(fin_c.all()) and bool(out["finite"].all())
    pos_g = out["pos"].cpu()
    extent = float(torch.cat([pos_c.reshape(-1, 3), pos_g.reshape(-1, 3)]).norm(dim=1).max())
    R = 2.0 * extent
    dR = R / 480
    sc = E.structure_set_stats(list(pos_c), R, dR)
    sg = E.structure_set_stats(list(pos_g), R, dR)
    cos = E.cos_similarity(sc["rdf"], sg["rdf"])
    wd = E.wasserstein(sc["rdf"], sg["rdf"])
    p_d = ks_2samp(sc["dists"], sg["dists"]).pvalue
    p_a = ks_2samp(sc["angles"], sg["angles"]).pvalue
    p_b = ks_2samp(sc["bonds"], sg["bonds"]).pvalue
    print(f"{precision}: RDF cos-sim {cos:.4f} wasserstein {wd:.3e} KS p: dist {p_d:.3f} angle {p_a:.3f} bond {p_b:.3f}")
    assert cos >= 0.99, cos
    assert min(p_d, p_a, p_b) > 0.01, (p_d, p_a, p_b)


def test_sampled_distributions_at_cfg3_shape(lib):
    """The same check on cfg3's shape -- 192-atom fully connected structures, EGNN L=6 d=128 (the sampling bench's net) --
    with a short schedule (T = 40) so that the CPU oracle finishes 12 structures in about a minute.  191 distances from
    atom 0 per structure: the local-RDF curves and the distance distributions of the two samplers must agree."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.sampler import sample_structures
    torch.manual_seed(1)
    L, d, n, T, S_cpu, S_gpu = 6, 128, 192, 40, 12, 192
    sizes = egnn_ref.net_sizes(L, d)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), L)
    sch = D.Schedule(0.05, 2.0, T)
    gen = torch.Generator().manual_seed(21)
    per = [LP.make_noises(n, T, 2, gen) for _ in range(S_cpu)]
    cat = {k: torch.cat([p[k] for p in per], dim=1 if k.startswith("step") else 0) for k in per[0]}
    with torch.no_grad():
        pos_c, _, fin_c = LP.sample_structures_batched(sd, L, sch, n, S_cpu, cat)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    net.precision = "bf16x3"
    dp = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=T)
    g = torch.Generator(device=DEV)
    g.manual_seed(22)
    out = sample_structures(net, dp, n, S_gpu, generator=g)
    assert bool(fin_c.all()) and bool(out["finite"].all())
    pos_g = out["pos"].cpu()
    extent = float(torch.cat([pos_c.reshape(-1, 3), pos_g.reshape(-1, 3)]).norm(dim=1).max())
    R = 2.0 * extent
    dR = R / 240
    sc = E.structure_set_stats(list(pos_c), R, dR)
    sg = E.structure_set_stats(list(pos_g), R, dR)
    cos = E.cos_similarity(sc["rdf"], sg["rdf"])
    p_d = ks_2samp(sc["dists"], sg["dists"]).pvalue
    print(f"cfg3 shape (192 atoms, L=6, d=128, T={T}): RDF cos-sim {cos:.4f} KS p dist {p_d:.3f}")
    assert cos >= 0.99, cos
    assert p_d > 0.01, p_d
