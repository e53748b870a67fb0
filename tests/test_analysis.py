"""Structure-analysis kernels against the reference's evaluation functions.

Golden 'eval_kat' holds outputs of the UNMODIFIED evaluate_RDF.RDF, CN2_evaluate.calculate_* and
evaluate_rmsd_for_pos_generate.kabsch_torch (oracle/make_eval_golden.py).  Bars: histogram counts, neighbour counts and
indices, distances: bit-exact; float64 RDF curves 1e-12 relative; angles 2e-3 degrees (fp32 acos, 1-2 ulp between libm and
CUDA); Kabsch rotation 1e-4 absolute, RMSD 1e-5 relative (fp32 SVD in the reference, float64 Jacobi here).
"""
import os

import numpy as np
import pytest
import torch

from oracle import eval_ref as E

torch.set_num_threads(4)


def load(golden_dir):
    return np.load(os.path.join(golden_dir, "eval_kat.npz"), allow_pickle=False)


def relerr(a, b):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def random_structures(S, n, seed, scale=2.5):
    g = torch.Generator().manual_seed(seed)
    return scale * torch.randn(S, n, 3, generator=g)


# ---------------------------------------------------------------- CPU: the oracle against the reference's outputs
def test_oracle_rdf_counts_and_edges(golden_dir):
    z = load(golden_dir)
    for k in ("0", "1", "2", "_edge"):
        R, dR = (z["par" + k] if k != "_edge" else (5.0, 0.01))
        pos = torch.from_numpy(z["pos" + k])
        assert np.array_equal(E.rdf(pos, R=float(R), dR=float(dR)), z["rdf" + k])
        assert np.array_equal(E.rdf_counts(pos, float(R), float(dR)), z["cnt" + k])
    # the adversarial cloud: 42 distances on / next to fp32 edge images; those exactly on an edge are in no bin
    assert int(z["cnt_edge"].sum()) == 26
    r, lo, hi = E.rdf_edges(5.0, 0.01)
    # in float64 r_k + dR != r_{k+1} for many k (the kernel tests neighbouring bins instead of assuming a partition);
    # their fp32 images happen to coincide for the default R, dR
    assert np.count_nonzero(r[1:] != (r + 0.01)[:-1]) > 0 and np.array_equal(lo[1:], hi[:-1])


def test_oracle_kabsch(golden_dir):
    z = load(golden_dir)
    for k in range(3):
        R, t, r = E.kabsch(z[f"kab_P{k}"], z[f"kab_Q{k}"])
        assert relerr(R, z[f"kab_R{k}"]) < 1e-5 and relerr(r, z[f"kab_rmsd{k}"]) < 1e-5
        assert np.allclose(t.numpy(), z[f"kab_t{k}"], atol=1e-6)
    # mirrored copy: the reference's column flip gives det +1 but a large RMSD (the quirk the kernels keep)
    assert float(z["kab_det1"]) > 0.99 and float(z["kab_rmsd1"]) > 10 * float(z["kab_rmsd0"])


# ---------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_rdf_kernel_golden(lib, golden_dir):
    from diffusion_model_b200 import evaluate as EV
    z = load(golden_dir)
    for k in ("0", "1", "2", "_edge"):
        R, dR = (z["par" + k] if k != "_edge" else (5.0, 0.01))
        pos = torch.from_numpy(z["pos" + k]).cuda()
        rdf, counts, dists = EV.RDF_batch([pos], R=float(R), dR=float(dR), return_counts=True, return_dists=True)
        assert np.array_equal(counts[0].cpu().numpy(), z["cnt" + k]), k
        want_d = torch.stack(E.length_from_exO(z["pos" + k]))
        assert torch.equal(dists[1:].cpu(), want_d), k  # torch.norm's fp32 arithmetic
        assert relerr(rdf[0], z["rdf" + k]) <= 1e-12, (k, relerr(rdf[0], z["rdf" + k]))
        got = EV.RDF(pos, R=float(R), dR=float(dR))
        assert isinstance(got, np.ndarray) and got.dtype == np.float64 and got.shape == z["rdf" + k].shape
        if k != "_edge":
            assert relerr(EV.RDF(pos, R=float(R), dR=float(dR), Normalize=True), z["rdfn" + k]) <= 1e-12
    ll = EV.length_from_exO(torch.from_numpy(z["pos0"]).cuda())
    assert len(ll) == 23 and all(float(a) == float(b) for a, b in zip(ll, E.length_from_exO(z["pos0"])))
    a, b = z["rdf0"], z["rdf1"][: len(z["rdf0"])]
    assert EV.cos_similarity(a, b) == float(z["cos01"])


@pytest.mark.gpu
def test_rdf_kernel_batches_and_edge_cases(lib):
    from diffusion_model_b200 import evaluate as EV
    pos = random_structures(48, 192, 5)
    rdf, counts = EV.RDF_batch(pos.cuda(), return_counts=True)
    for s in range(0, 48, 5):
        assert np.array_equal(counts[s].cpu().numpy(), E.rdf_counts(pos[s])), s
        assert relerr(rdf[s], E.rdf(pos[s])) <= 1e-12, s
    # ragged list, a one-atom structure (no distances), sigma / R / dR variants, empty batch
    rag = [pos[0, :17], pos[1, :1], pos[2, :100]]
    rdf, counts = EV.RDF_batch([p.cuda() for p in rag], sigma=3, R=8.0, dR=0.05, return_counts=True)
    for s, p in enumerate(rag):
        assert np.array_equal(counts[s].cpu().numpy(), E.rdf_counts(p, 8.0, 0.05)), s
        if len(p) > 1:
            assert relerr(rdf[s], E.rdf(p, sigma=3, R=8.0, dR=0.05)) <= 1e-12
    assert int(counts[1].sum()) == 0 and float(rdf[1].abs().max()) == 0.0
    assert EV.RDF_batch(torch.zeros(0, 4, 3).cuda()).shape == (0, 500)
    # reruns are bit-identical (integer atomics + fixed-order float64)
    r2 = EV.RDF_batch(pos.cuda())
    assert torch.equal(r2, EV.RDF_batch(pos.cuda()))


@pytest.mark.gpu
def test_rdf_kernel_full_size_counts(lib):
    """cfg3 size: 4096 structures x 192 atoms; the histogram against an independent device-side count (broadcast
    comparisons with the same fp32 edges), bit-exact; total = number of distances inside some bin"""
    from diffusion_model_b200 import evaluate as EV
    pos = random_structures(4096, 192, 6).cuda()
    rdf, counts, dists = EV.RDF_batch(pos, return_counts=True, return_dists=True)
    _, lo, hi = E.rdf_edges(5.0, 0.01)
    lo, hi = torch.from_numpy(lo).cuda(), torch.from_numpy(hi).cuda()
    d = (pos[:, 1:] - pos[:, :1]).norm(dim=2)
    assert relerr(dists.view(4096, 192)[:, 1:], d) < 1e-6
    d = dists.view(4096, 192)[:, 1:]
    for s0 in range(0, 4096, 512):
        dd = d[s0:s0 + 512].unsqueeze(2)
        want = ((lo < dd) & (dd < hi)).sum(dim=1).to(torch.int32)
        assert torch.equal(counts[s0:s0 + 512], want), s0
    assert torch.isfinite(rdf).all()


@pytest.mark.gpu
def test_cn2_kernel(lib, golden_dir):
    from diffusion_model_b200 import evaluate as EV
    z = load(golden_dir)
    for k in range(3):
        tri = torch.from_numpy(z[f"pos{k}"][:3]).cuda()
        ang, b1, b2 = z[f"cn2_{k}"]
        assert abs(EV.calculate_angle_for_CN2(tri) - ang) <= 2e-3
        assert EV.calculate_bond_length_for_CN2(tri) == (float(b1), float(b2))
    # site selection by radius (evaluate_Si-O-Si.py:28-39) on clouds dense enough to have 0..4 neighbours
    pos = random_structures(200, 24, 8, scale=1.6)
    out = EV.cn2_sites(pos.cuda(), radius=2.0)
    cnt, idx = out["count"].cpu(), out["idx"].cpu()
    seen = set()
    for s in range(200):
        nb, ang, bonds = E.cn2_site(pos[s], 2.0)
        seen.add(min(len(nb), 3))
        assert int(cnt[s]) == len(nb), s
        assert idx[s].tolist() == (nb + [-1, -1])[:2], s
        if len(nb) == 2:
            assert abs(float(out["angle"][s]) - ang) <= 2e-3 or abs(ang - 90) > 89.9, (s, ang)
            assert tuple(out["bonds"][s].cpu().tolist()) == tuple(float(np.float32(b)) for b in bonds)
    assert seen == {0, 1, 2, 3}
    # nearest-two rule (used for untrained samplers, oracle/eval_ref.two_nearest)
    near = EV.cn2_sites(pos.cuda(), radius=None)
    for s in range(0, 200, 7):
        assert [0] + near["idx"][s].tolist() == E.two_nearest(pos[s]), s


@pytest.mark.gpu
def test_kabsch_kernel(lib, golden_dir):
    from diffusion_model_b200 import evaluate as EV
    z = load(golden_dir)
    for k in range(3):
        R, t, r = EV.kabsch_torch(torch.from_numpy(z[f"kab_P{k}"]).cuda(), torch.from_numpy(z[f"kab_Q{k}"]).cuda())
        assert float((R.cpu() - torch.from_numpy(z[f"kab_R{k}"])).abs().max()) <= 1e-4, k
        assert np.allclose(t.cpu().numpy(), z[f"kab_t{k}"], atol=1e-5)
        assert abs(float(r) - float(z[f"kab_rmsd{k}"])) <= 1e-5 * float(z[f"kab_rmsd{k}"]), k
    # batch of rotated + noisy copies, every third one mirrored, against the oracle restatement
    g = torch.Generator().manual_seed(3)
    P = random_structures(30, 192, 9)
    Q = torch.empty_like(P)
    for s in range(30):
        A = torch.linalg.qr(torch.randn(3, 3, generator=g))[0]
        if (torch.det(A) < 0) != (s % 3 == 0):
            A[:, 0] *= -1
        Q[s] = P[s] @ A.T + 0.05 * torch.randn(192, 3, generator=g) + torch.randn(3, generator=g)
    R, t, r = EV.kabsch_batch(P.cuda(), Q.cuda())
    for s in range(30):
        Ro, to, ro = E.kabsch(P[s], Q[s])
        assert float((R[s].cpu() - Ro).abs().max()) <= 1e-4 and abs(float(r[s]) - float(ro)) <= 1e-5 * float(ro), s
        assert float((t[s].cpu() - to).abs().max()) <= 1e-5
    assert torch.equal(r, EV.kabsch_batch(P.cuda(), Q.cuda())[2])
