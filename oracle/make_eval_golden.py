"""Pins oracle/eval_ref.py against the UNMODIFIED reference evaluation functions (run in the build container only).

Imports /root/reference/evaluate_RDF.py and CN2_evaluate.py with their plotting / logging imports stubbed
(matplotlib, wandb, PIL, split_to_train_and_test are not on the path of the functions used), evaluates them on seeded
point clouds, asserts the restatement reproduces them and freezes the reference's outputs in
tests/golden/eval_kat.npz.   python -m oracle.make_eval_golden
"""
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def import_reference():
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.gridspec", "wandb", "PIL", "PIL.Image"):
        if name not in sys.modules:
            _stub(name)
    sys.modules["matplotlib.gridspec"].GridSpec = object
    sys.modules["matplotlib.gridspec"].GridSpecFromSubplotSpec = object
    sys.modules["PIL"].Image = sys.modules["PIL.Image"]
    _stub("split_to_train_and_test", SetUpData=object)
    sys.path.insert(0, REF)
    import evaluate_RDF
    import CN2_evaluate
    import evaluate_rmsd_for_pos_generate as rmsd_mod
    return evaluate_RDF, CN2_evaluate, rmsd_mod


def main():
    from oracle import eval_ref as E
    ref_rdf, ref_cn2, ref_rmsd = import_reference()
    gen = torch.Generator().manual_seed(2024)
    out = {}
    for k, (n, scale, R, dR) in enumerate([(24, 2.0, 5.0, 0.01), (64, 3.0, 5.0, 0.01), (24, 30.0, 120.0, 0.25)]):
        pos = scale * torch.randn(n, 3, generator=gen)
        want = ref_rdf.RDF(pos, R=R, dR=dR)
        want_n = ref_rdf.RDF(pos, R=R, dR=dR, Normalize=True)
        got = E.rdf(pos, R=R, dR=dR)
        assert np.array_equal(np.asarray(want), got), k
        assert np.array_equal(np.asarray(want_n), E.rdf(pos, R=R, dR=dR, Normalize=True)), k
        out[f"pos{k}"], out[f"rdf{k}"], out[f"rdfn{k}"] = pos.numpy(), np.asarray(want), np.asarray(want_n)
        out[f"par{k}"] = np.array([R, dR])
        tri = pos[:3]
        ang, (b1, b2) = ref_cn2.calculate_angle_for_CN2(tri), ref_cn2.calculate_bond_length_for_CN2(tri)
        assert ang == E.angle_cn2(tri) and (b1, b2) == E.bond_lengths_cn2(tri)
        out[f"cn2_{k}"] = np.array([ang, b1, b2])
        out[f"cnt{k}"] = E.rdf_counts(pos, R=R, dR=dR)
    # adversarial cloud: distances sitting EXACTLY on the fp32 images of bin edges and one ulp either side -- the
    # reference's fp32 comparison (see eval_ref.rdf_edges) decides these, a float64 comparison would not
    _, lo, hi = E.rdf_edges(5.0, 0.01)
    ds = []
    for k in (0, 4, 17, 99, 250, 498, 499):
        for v in (lo[k], hi[k]):
            ds += [v, np.nextafter(v, np.float32(0)), np.nextafter(v, np.float32(10))]
    pos = torch.zeros(len(ds) + 1, 3)
    pos[1:, 0] = torch.tensor(np.array(ds, dtype=np.float32)) * torch.tensor([1.0, -1.0] * (len(ds) // 2))
    want = ref_rdf.RDF(pos, R=5.0, dR=0.01)
    assert np.array_equal(np.asarray(want), E.rdf(pos, R=5.0, dR=0.01)), "edge-case cloud"
    out["pos_edge"], out["rdf_edge"], out["cnt_edge"] = pos.numpy(), np.asarray(want), E.rdf_counts(pos, 5.0, 0.01)
    # Kabsch alignment (evaluate_rmsd_for_pos_generate.py:11-51): a rotated + noisy copy and a MIRRORED copy (det < 0 branch)
    for k, (n, mirror) in enumerate([(24, False), (24, True), (192, False), (3, True)]):
        P = 3.0 * torch.randn(n, 3, generator=gen)
        A_ = torch.linalg.qr(torch.randn(3, 3, generator=gen))[0]
        if (torch.det(A_) < 0) != mirror:
            A_[:, 0] *= -1.0
        Q = P @ A_.T + 0.1 * torch.randn(n, 3, generator=gen) + torch.tensor([1.0, -2.0, 0.5])
        Rr, tr, rr = ref_rmsd.kabsch_torch(P, Q)
        Ro, to, ro = E.kabsch(P, Q)
        assert torch.equal(Rr, Ro) and torch.equal(tr, to) and torch.equal(rr, ro), k
        out[f"kab_P{k}"], out[f"kab_Q{k}"] = P.numpy(), Q.numpy()
        out[f"kab_R{k}"], out[f"kab_t{k}"], out[f"kab_rmsd{k}"] = Rr.numpy(), tr.numpy(), np.array(rr.item())
        out[f"kab_det{k}"] = np.array(torch.det(Rr).item())
    a, b = out["rdf0"], out["rdf1"][: len(out["rdf0"])]
    assert ref_rdf.cos_similarity(a, b) == E.cos_similarity(a, b)
    assert ref_rdf.calculate_wasserstein_distance(a, b) == E.wasserstein(a, b)
    out["cos01"], out["wd01"] = np.array(ref_rdf.cos_similarity(a, b)), np.array(ref_rdf.calculate_wasserstein_distance(a, b))
    np.savez_compressed(os.path.join(GOLD, "eval_kat.npz"), **out)
    print("eval_kat.npz written;", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
