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
    return evaluate_RDF, CN2_evaluate


def main():
    from oracle import eval_ref as E
    ref_rdf, ref_cn2 = import_reference()
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
    a, b = out["rdf0"], out["rdf1"][: len(out["rdf0"])]
    assert ref_rdf.cos_similarity(a, b) == E.cos_similarity(a, b)
    assert ref_rdf.calculate_wasserstein_distance(a, b) == E.wasserstein(a, b)
    out["cos01"], out["wd01"] = np.array(ref_rdf.cos_similarity(a, b)), np.array(ref_rdf.calculate_wasserstein_distance(a, b))
    np.savez_compressed(os.path.join(GOLD, "eval_kat.npz"), **out)
    print("eval_kat.npz written;", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
