"""CPU tests: the oracle restatement against the golden vectors frozen from the UNMODIFIED reference
(oracle/make_golden.py), and the C-ABI surface (library loads, exports every symbol the header declares)."""
import os
import re

import numpy as np
import torch

from oracle import diffusion_ref as D
from oracle import egnn_ref, graphs_ref as G, loops_ref as LP

torch.set_num_threads(4)


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def state_dict_of(z):
    return {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w:")}


def sizes_of(z):
    return {str(k): int(v) for k, v in zip(z["size_keys"], z["sizes"])}


def relerr(a, b):
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def test_schedule_kat(golden_dir):
    z = load(golden_dir, "diffusion_kat")
    sch = D.Schedule(float(z["s"]), float(z["power"]), int(z["T"]))
    assert np.array_equal(sch.alpha.numpy(), z["alpha"])
    assert np.array_equal(sch.sigma.numpy(), z["sigma"])
    # SURVEY 3.4 known answers
    assert abs(float(sch.alpha[500]) - 0.56249851) < 1e-7
    assert abs(float(sch.sigma[1000]) - 1.0) < 1e-7


def test_remove_mean_kat(golden_dir):
    z = load(golden_dir, "diffusion_kat")
    out = D.remove_mean(torch.from_numpy(z["rm_in"]).clone(), torch.from_numpy(z["rm_batch"]))
    assert np.array_equal(out.numpy(), z["rm_out"])
    assert np.array_equal(D.remove_mean(torch.from_numpy(z["rm_in"]).clone()).numpy(), z["rm_out_global"])


def test_reverse_step_and_mu(golden_dir):
    z = load(golden_dir, "diffusion_kat")
    sch = D.Schedule(float(z["s"]), float(z["power"]), int(z["T"]))
    zz, eps = torch.from_numpy(z["z"]), torch.from_numpy(z["eps"])
    zh, eh = torch.from_numpy(z["zh"]), torch.from_numpy(z["eh"])
    for t in z["ts"].tolist():
        assert np.array_equal(D.calculate_mu(sch, zz, eps, t).numpy(), z[f"t{t}_mu"])
        out = D.reverse_diffuse_one_step(sch, zz, eps, t, torch.from_numpy(z[f"t{t}_noise"]).clone(), "pos")
        assert np.array_equal(out.numpy(), z[f"t{t}_z"])
        outh = D.reverse_diffuse_one_step(sch, zh, eh, t, torch.from_numpy(z[f"t{t}_noise_h"]).clone(), "h")
        assert np.array_equal(outh.numpy(), z[f"t{t}_zh"])
        zt, nz = D.diffuse_zero_to_t(sch, zz, t, torch.from_numpy(z[f"t{t}_raw"]).clone(), "pos")
        assert np.array_equal(zt.numpy(), z[f"t{t}_zt"]) and np.array_equal(nz.numpy(), z[f"t{t}_nz"])
        assert np.array_equal(D.calculate_mu_xonly(sch, zz, eps, t).numpy(), z[f"t{t}_mu_xonly"])


def test_edge_builders(golden_dir):
    z = load(golden_dir, "edges")
    for n in (1, 2, 5, 24):
        assert np.array_equal(G.fc_edges(n), z[f"fc_{n}"].reshape(2, -1))
    for k in ("dense", "sparse", "lattice"):
        assert np.array_equal(G.cutoff_edges(z[f"open_{k}_pos"], 2.0, None), z[f"open_{k}_edges"])
    e = G.cutoff_edges_batch(z["per_pos"], z["per_graph_ptr"], 5.0, z["per_boxes"])
    assert np.array_equal(e, z["per_edges"])


def test_egnn_layers_against_reference(golden_dir):
    for name in ("cfg1", "ragged"):
        z = load(golden_dir, "egnn_" + name)
        sd, sz = state_dict_of(z), sizes_of(z)
        h, x, ei = torch.from_numpy(z["h"]), torch.from_numpy(z["x"]), torch.from_numpy(z["edge_index"])
        with torch.no_grad():
            _, _, layers = egnn_ref.egnn_forward(sd, sz["L"], ei, h, x, "call", return_layers=True)
            hg, xg = egnn_ref.egnn_forward(sd, sz["L"], ei, h, x, "graph", batch=torch.from_numpy(z["batch"]))
        for l, (a, c) in enumerate(layers):
            assert relerr(a, z[f"h_l{l}"]) < 5e-6 and relerr(c, z[f"x_l{l}"]) < 5e-6
        assert relerr(hg, z["h_graph"]) < 5e-6 and relerr(xg, z["x_graph"]) < 5e-6


def test_loss_against_reference(golden_dir):
    z = load(golden_dir, "loss_cfg1")
    sd = state_dict_of(z)
    sch = D.Schedule(1e-5, 2.0, 1000)
    egnn = lambda e, h, x: egnn_ref.egnn_forward(sd, 4, e, h, x)
    with torch.no_grad():
        loss, aux = LP.train_loss(egnn, sch, torch.from_numpy(z["pos"]), torch.from_numpy(z["x"]),
                                  torch.from_numpy(z["edge_index"]), torch.from_numpy(z["batch"]),
                                  torch.from_numpy(z["t"]), torch.from_numpy(z["noise_pos"]), torch.from_numpy(z["noise_h"]))
    assert abs(float(loss) - float(z["loss"])) < 1e-5 * abs(float(z["loss"]))
    assert np.array_equal(aux["pos_t"].numpy(), z["pos_t"])


def test_abi_exports_every_declared_symbol(lib):
    from diffusion_model_b200 import _lib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "egnn_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(egnn_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.SIGNATURES), (declared ^ set(_lib.SIGNATURES))
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.egnn_b200_version() == 1
    net = _lib.NetDesc(6, 3, 128, 128, 128, 128)
    import ctypes as C
    assert lib.egnn_param_count(C.byref(net)) == 315678  # cfg 2 network
    net = _lib.NetDesc(4, 3, 64, 64, 64, 64)
    assert lib.egnn_param_count(C.byref(net)) == 56084


def test_product_refuses_cpu_tensors(lib):
    import pytest
    from diffusion_model_b200 import EgnnError
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    net = EquivariantGNN(**egnn_ref.net_sizes(1, 8))
    ei = torch.from_numpy(G.fc_edges(4))
    with pytest.raises(EgnnError):
        net(ei, torch.zeros(4, 3), torch.zeros(4, 3))


def test_state_dict_keys_match_reference(golden_dir):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    z = load(golden_dir, "egnn_cfg1")
    sd, sz = state_dict_of(z), sizes_of(z)
    torch.manual_seed(2024)
    net = EquivariantGNN(**sz)
    mine = net.state_dict()
    assert list(mine.keys()) == list(sd.keys())
    # same construction order under the same seed -> identical initial weights as the reference module
    assert all(torch.equal(mine[k], sd[k]) for k in sd)


def test_eval_functions_against_reference(golden_dir):
    """RDF / CN2 angle / bond-length restatement (oracle/eval_ref.py) vs outputs of the unmodified reference functions"""
    from oracle import eval_ref as E
    z = load(golden_dir, "eval_kat")
    for k in range(3):
        R, dR = (float(v) for v in z[f"par{k}"])
        pos = torch.from_numpy(z[f"pos{k}"])
        assert np.array_equal(E.rdf(pos, R=R, dR=dR), z[f"rdf{k}"])
        assert np.array_equal(E.rdf(pos, R=R, dR=dR, Normalize=True), z[f"rdfn{k}"])
        ang, b1, b2 = z[f"cn2_{k}"]
        assert E.angle_cn2(pos[:3]) == ang and E.bond_lengths_cn2(pos[:3]) == (b1, b2)
    a, b = z["rdf0"], z["rdf1"][: len(z["rdf0"])]
    assert E.cos_similarity(a, b) == float(z["cos01"]) and E.wasserstein(a, b) == float(z["wd01"])


def test_batched_oracle_sampler_equals_per_structure_loop():
    """the batched CPU sampler used by the distribution test is the per-structure loop of generate() run S times"""
    torch.manual_seed(0)
    sizes = egnn_ref.net_sizes(2, 32)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), 2)
    n, S, T = 12, 3, 40
    sch = D.Schedule(0.05, 2.0, T)
    gen = torch.Generator().manual_seed(1)
    per = [LP.make_noises(n, T, 2, gen) for _ in range(S)]
    cat = {k: torch.cat([p[k] for p in per], dim=1 if k.startswith("step") else 0) for k in per[0]}
    ei = torch.as_tensor(G.fc_edges(n))
    with torch.no_grad():
        pos, hf, fin = LP.sample_structures_batched(sd, 2, sch, n, S, cat)
        for k in range(S):
            p1, h1, _, f1 = LP.sample_structure(lambda e, h, x: egnn_ref.egnn_forward(sd, 2, e, h, x), sch, ei, n, per[k])
            assert bool(fin[k]) == f1
            assert float((pos[k] - p1).abs().max()) <= 1e-4 * float(p1.abs().max())
            assert float((hf[k] - h1).abs().max()) <= 1e-4 * max(1.0, float(h1.abs().max()))


def test_radam_schedulefree_restatement_invariants():
    """oracle/radam_sf_ref.py (parity unpinned: the schedulefree package is absent): invariants of the published algorithm --
    the first steps (rho_t <= 4) are silent, x_t is the c_t-weighted running average of the z sequence with
    y_t = beta1 x_t + (1 - beta1) z_t, eval()/train() are inverse maps, and it minimises a quadratic."""
    from oracle.radam_sf_ref import RAdamScheduleFreeRef
    torch.manual_seed(0)
    target = torch.randn(32)
    p = torch.zeros(32)
    opt = RAdamScheduleFreeRef([p], lr=0.05)
    opt.train()
    x_avg = None
    for it in range(400):
        p.grad = p - target
        z_before = None if p not in opt.state else opt.state[p]["z"].clone()
        ws_before = opt.g["weight_sum"]
        opt.step()
        if it < 4:
            assert torch.equal(p, torch.zeros(32)) and opt.g["scheduled_lr"] == 0.0
        z = opt.state[p]["z"]
        w = opt.g["weight_sum"] - ws_before
        c = w / opt.g["weight_sum"] if opt.g["weight_sum"] > 0 else 0.0
        x_avg = z.clone() if x_avg is None else (1 - c) * x_avg + c * z
        if opt.g["weight_sum"] > 0:
            assert torch.allclose(p, 0.9 * x_avg + 0.1 * z, atol=1e-5), it
    opt.eval()
    assert torch.allclose(p, x_avg, atol=1e-5)
    assert float((p - target).abs().max()) < 0.15 < float(target.abs().max())  # from 2.1 down, 400 Adam-sized steps
    opt.train(); opt.eval()
    assert torch.allclose(p, x_avg, atol=1e-5)


def test_default_precision_resolution():
    """host logic of the drop-in: swapping the imports alone selects the tcgen05 bf16x3 mode wherever the tensor-core
    kernels cover the shape (same 1e-4 bar as the fp32 mode), and the CUDA-core fp32 mode otherwise -- never 'bf16'."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EGCL, EquivariantGNN
    for d in (64, 128, 256):
        assert EquivariantGNN(**egnn_ref.net_sizes(2, d)).resolved_precision() == "bf16x3"
    assert EquivariantGNN(**egnn_ref.net_sizes(1, 96)).resolved_precision() == "fp32"            # no tile for d = 96
    assert EquivariantGNN(**egnn_ref.net_sizes(1, 64, H=36)).resolved_precision() == "fp32"      # K0 = 73 > 16, narrow
    # the reference's parameters.yaml widths: the tile-GEMM forward (csrc/tc_wide.cu)
    yaml = dict(L=1, m_input=73, m_hidden=1024, m_output=256, x_input=73, x_hidden=1024, x_output=1, h_input=36 + 256,
                h_hidden=1024, h_output=36)
    assert EquivariantGNN(**yaml).resolved_precision() == "bf16x3"
    assert EquivariantGNN(**egnn_ref.net_sizes(1, 512, H=70)).resolved_precision() == "fp32"     # K0 = 141 > 128
    sz = egnn_ref.net_sizes(1, 128)
    sz.pop("L")
    layer = EGCL(**sz)
    assert layer.precision == "auto" and layer.resolved_precision() == "bf16x3"
    layer.precision = "fp32"
    assert layer.resolved_precision() == "fp32"


def test_workspace_and_image_sizes_of_the_tile_gemm_path(lib):
    """Host-side layout arithmetic of csrc/common.cuh (no GPU needed): the reference's parameters.yaml widths get tile-GEMM
    weight images and forward / backward scratch, the fused-kernel networks of BASELINE.json do not pay for them."""
    import ctypes as C
    from diffusion_model_b200 import _lib
    yaml = _lib.NetDesc(5, 36, 1024, 256, 1024, 1024)
    P = lib.egnn_param_count(C.byref(yaml))
    assert P == 9008830                                    # 5 layers x 1 801 766 (main.py:112-122 with the yaml's sizes)
    packed = lib.egnn_packed_count(C.byref(yaml))
    # forward images m1 8 + m2 16 + x1 8 + x2 64, dgrad images 16 + 64 + 32, node 20 + 16 + 4 + 32 blocks of 64 KB per layer
    blocks = 8 + 16 + 8 + 64 + 16 + 64 + 32 + 4 * 5 + 16 + 4 * 1 + 2 * 16
    assert packed % 4 == 0 and packed >= 5 * blocks * 16384
    assert packed < 5 * (blocks * 16384 + 2 * 1801766)     # + the K-major fp32 transposes of the CUDA-core path
    E, N, G = 9120, 96, 1
    f0 = lib.egnn_forward_workspace_floats(C.byref(yaml), N, 0, G)
    f1 = lib.egnn_forward_workspace_floats(C.byref(yaml), N, E, G)
    mt = (E + 127) // 128
    # U image (2 slices) + A1 image (16 slices) of 32 KB per 128-edge tile, gated messages [E x 256] fp32, 4 phi parts
    want = mt * 18 * 32768 + E * 256 * 4 + 4 * E * 4
    assert (f1 - f0) * 4 >= want and (f1 - f0) * 4 < want + (2 * E + 8) * 4 * 300 + (1 << 20)
    b1 = lib.egnn_backward_workspace_floats(C.byref(yaml), N, E, G)
    img = mt * 16 * 32768                                  # one [E x 1024] hi | lo image
    assert b1 * 4 >= 8 * img + 2 * mt * 128 * 1024 * 4     # a1, a1T, g2, g2T, g1m, g1x, g1T (+u) images; SiLU' rows and row sums
    # ... plus the partial-row buffer, sized for the fp32 path too (up to 148 edge rows + node rows of one layer's parameters)
    assert b1 * 4 < 16 * img + 2 * 148 * 1801766 * 4
    # cfg2's network: fused kernels only -- no tile-GEMM images, no tile-GEMM scratch
    cfg2 = _lib.NetDesc(6, 3, 128, 128, 128, 128)
    assert lib.egnn_packed_count(C.byref(cfg2)) < 6 * 200000
    assert lib.egnn_forward_workspace_floats(C.byref(cfg2), 12288, 417000, 64) * 4 < 64 << 20
