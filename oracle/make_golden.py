"""TEST INFRASTRUCTURE ONLY.  Run in the BUILD container (needs /root/reference, read-only):

    python -m oracle.make_golden            # validates the restatement against the reference, writes tests/golden/

1. imports the reference's EquivariantGraphNeuralNetwork.py, diffusion_x_h.py and E3diffusion_new.py UNMODIFIED
   (the first through oracle/pyg_stub.py, because torch_geometric is absent here),
2. asserts that oracle/{egnn_ref,diffusion_ref,graphs_ref,loops_ref}.py reproduce the reference on seeded inputs,
3. freezes the REFERENCE's outputs as small fixtures under tests/golden/ (these travel to the GPU box;
   /root/reference does not).
Nothing in tests/, smoke() or bench.py reads /root/reference at run time.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF = os.environ.get("EGNN_REFERENCE_DIR", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

from . import pyg_stub, egnn_ref, diffusion_ref as D, graphs_ref as G, loops_ref as LP  # noqa: E402


def load_reference():
    pyg_stub.install()
    sys.path.insert(0, REF)
    import importlib
    egnn_mod = importlib.import_module("EquivariantGraphNeuralNetwork")
    dxh = importlib.import_module("diffusion_x_h")
    dnew = importlib.import_module("E3diffusion_new")
    return egnn_mod, dxh, dnew


def relerr(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


NETS = {
    # name: (sizes, graph spec)
    "cfg1": (egnn_ref.net_sizes(4, 64), dict(num_graphs=8, n_atoms=24, edges="fc", seed=11)),
    "cfg2": (egnn_ref.net_sizes(6, 128), dict(num_graphs=2, n_atoms=192, edges="cutoff_periodic", seed=12)),
    "cfg4": (egnn_ref.net_sizes(2, 256), dict(num_graphs=1, n_atoms=96, edges="cutoff_periodic", seed=13)),
    "ragged": (dict(L=2, m_input=2 * 9 + 1, m_hidden=96, m_output=64, x_input=2 * 9 + 1, x_hidden=80, x_output=1,
                    h_input=9 + 64, h_hidden=48, h_output=9), dict(num_graphs=3, n_atoms=17, edges="cutoff_open", seed=14)),
}


def golden_egnn(egnn_mod):
    for name, (sizes, gs) in NETS.items():
        torch.manual_seed(2024)
        ref = egnn_mod.EquivariantGNN(**sizes)
        sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
        torch.manual_seed(2024)
        sd2 = egnn_ref.init_state_dict(**sizes)
        assert all(torch.equal(sd[k], sd2[k]) for k in sd), "init order differs from the reference"
        b = G.make_batch(**gs)
        N = b["pos"].shape[0]
        H = sizes["h_output"]
        gen = torch.Generator().manual_seed(77)
        h = torch.randn(N, H, generator=gen)
        x = torch.from_numpy(b["pos"]) + 0.1 * torch.randn(N, 3, generator=gen)
        ei = torch.from_numpy(b["edge_index"])
        batch = torch.from_numpy(b["batch"])
        # reference, layer by layer (call-scope norm is what a batched reference call computes)
        hh, xx, layers = h, x, []
        with torch.no_grad():
            for l in range(sizes["L"]):
                hh, xx = ref.egcl_list[l](ei, hh, xx)
                layers.append((hh, xx))
            h_ref, x_ref = ref(ei, h, x)
            h_o, x_o, lay_o = egnn_ref.egnn_forward(sd, sizes["L"], ei, h, x, "call", return_layers=True)
        assert torch.equal(h_ref, layers[-1][0])
        e = max(relerr(h_o, h_ref), relerr(x_o, x_ref))
        assert e < 2e-6, (name, e)
        # graph-scope norm == reference run one graph at a time
        with torch.no_grad():
            h_g, x_g = egnn_ref.egnn_forward(sd, sizes["L"], ei, h, x, "graph", batch=batch)
            hs, xs = [], []
            for g in range(b["num_graphs"]):
                a, c = int(b["graph_ptr"][g]), int(b["graph_ptr"][g + 1])
                sel = (ei[0] >= a) & (ei[0] < c)
                hg, xg = ref(ei[:, sel] - a, h[a:c], x[a:c])
                hs.append(hg); xs.append(xg)
        e2 = max(relerr(h_g, torch.cat(hs)), relerr(x_g, torch.cat(xs)))
        assert e2 < 2e-6, (name, e2)
        # loss-like scalar + grads from the REFERENCE module (autograd)
        ref.zero_grad()
        gen2 = torch.Generator().manual_seed(5)
        th, tx = torch.randn(N, H, generator=gen2), torch.randn(N, 3, generator=gen2)
        hr, xr = ref(ei, h, x)
        loss = ((hr - th) ** 2).sum() + ((xr - x - tx) ** 2).sum()
        loss.backward()
        out = dict(h=h.numpy(), x=x.numpy(), edge_index=b["edge_index"], batch=b["batch"], target_h=th.numpy(),
                   target_x=tx.numpy(), loss=np.float64(loss.item()), h_graph=torch.cat(hs).numpy(),
                   x_graph=torch.cat(xs).numpy(), sizes=np.array(list(sizes.values()), np.int64),
                   size_keys=np.array(list(sizes.keys())))
        for l, (a, c) in enumerate(layers):
            out[f"h_l{l}"] = a.numpy(); out[f"x_l{l}"] = c.numpy()
        for k, v in sd.items():
            out["w:" + k] = v.numpy()
        for k, p in ref.named_parameters():
            out["g:" + k] = p.grad.numpy()
        np.savez_compressed(os.path.join(GOLD, f"egnn_{name}.npz"), **out)
        print(f"egnn_{name}: N={N} E={ei.shape[1]} restatement-vs-reference relerr call={e:.2e} graph={e2:.2e}")


def golden_diffusion(dxh, dnew):
    T, s, power = 1000, 1e-5, 2.0
    ref = dxh.E3DiffusionProcess(s=s, power=power, num_diffusion_timestep=T)
    refn = dnew.E3DiffusionProcess(s=s, power=power, num_diffusion_timestep=T)
    sch = D.Schedule(s, power, T)
    assert torch.equal(ref.alpha_schedule, sch.alpha) and torch.equal(ref.sigma_schedule, sch.sigma)
    assert torch.equal(refn.alpha_schedule, sch.alpha)
    out = dict(alpha=sch.alpha.numpy(), sigma=sch.sigma.numpy(), T=np.int64(T), s=np.float64(s), power=np.float64(power))
    # remove_mean KAT
    x = torch.arange(12, dtype=torch.float32).view(4, 3)
    b = torch.tensor([0, 0, 1, 1])
    r1 = dxh.remove_mean(x.clone(), b)
    assert torch.equal(r1, D.remove_mean(x.clone(), b))
    assert torch.equal(dxh.remove_mean(x.clone()), D.remove_mean(x.clone()))
    out["rm_in"], out["rm_batch"], out["rm_out"] = x.numpy(), b.numpy(), r1.numpy()
    out["rm_out_global"] = dxh.remove_mean(x.clone()).numpy()
    # mu / reverse step / diffuse with the reference's own RNG draws
    gen = torch.Generator().manual_seed(3)
    z, eps = torch.randn(24, 3, generator=gen), torch.randn(24, 3, generator=gen)
    zh, eh = torch.randn(24, 2, generator=gen), torch.randn(24, 2, generator=gen)
    ts = [1, 2, 17, 500, 999, 1000]
    out["ts"] = np.array(ts)
    out["z"], out["eps"], out["zh"], out["eh"] = z.numpy(), eps.numpy(), zh.numpy(), eh.numpy()
    for t in ts:
        mu = ref.calculate_mu(z, eps, t)
        assert torch.equal(mu, D.calculate_mu(sch, z, eps, t))
        torch.manual_seed(100 + t)
        zr = ref.reverse_diffuse_one_step(z, eps, t, mode="pos")
        torch.manual_seed(100 + t)
        noise = torch.zeros_like(z).normal_(mean=0, std=1)
        assert torch.equal(zr, D.reverse_diffuse_one_step(sch, z, eps, t, noise.clone(), "pos"))
        torch.manual_seed(200 + t)
        zhr = ref.reverse_diffuse_one_step(zh, eh, t, mode="h")
        torch.manual_seed(200 + t)
        noise_h = torch.zeros_like(zh).normal_(mean=0, std=1)
        assert torch.equal(zhr, D.reverse_diffuse_one_step(sch, zh, eh, t, noise_h.clone(), "h"))
        torch.manual_seed(300 + t)
        zt, nz = ref.diffuse_zero_to_t(z, t, mode="pos")
        torch.manual_seed(300 + t)
        raw = torch.zeros_like(z).normal_(mean=0, std=1)
        zt2, nz2 = D.diffuse_zero_to_t(sch, z, t, raw.clone(), "pos")
        assert torch.equal(zt, zt2) and torch.equal(nz, nz2)
        # x-only class
        mun = refn.calculate_mu(z, eps, t)
        assert torch.equal(mun, D.calculate_mu_xonly(sch, z, eps, t))
        torch.manual_seed(400 + t)
        pn = refn.reverse_diffuse_one_step(mun, t)
        torch.manual_seed(400 + t)
        rawn = torch.zeros_like(z).normal_(mean=0, std=1)
        assert torch.equal(pn, D.reverse_diffuse_one_step_xonly(sch, mun, t, rawn.clone()))
        a_ts, s2_ts, sig_t, std = sch.step_scalars(t)
        out[f"t{t}_scalars"] = np.array([a_ts, s2_ts, sig_t, std], np.float32)
        out[f"t{t}_mu"], out[f"t{t}_noise"], out[f"t{t}_z"] = mu.numpy(), noise.numpy(), zr.numpy()
        out[f"t{t}_noise_h"], out[f"t{t}_zh"] = noise_h.numpy(), zhr.numpy()
        out[f"t{t}_raw"], out[f"t{t}_zt"], out[f"t{t}_nz"] = raw.numpy(), zt.numpy(), nz.numpy()
        out[f"t{t}_mu_xonly"], out[f"t{t}_rawn"], out[f"t{t}_pos_xonly"] = mun.numpy(), rawn.numpy(), pn.numpy()
    # t/T column: train_epoch (long tensor / int) vs generate (python float -> fp32): identical for T=1000?
    tt = torch.arange(0, T + 1)
    col_a = (tt / T).to(torch.float32)
    col_b = torch.tensor([t / T for t in range(T + 1)], dtype=torch.float32)
    out["time_cols_equal"] = np.bool_(torch.equal(col_a, col_b))
    out["time_col"] = col_a.numpy()
    np.savez_compressed(os.path.join(GOLD, "diffusion_kat.npz"), **out)
    print("diffusion_kat: schedule/mu/reverse/diffuse bit-equal to reference; time columns equal:", bool(out["time_cols_equal"]))


def golden_edges():
    """Edge builders against the notebook's torch.norm double loop (something.ipynb:6340-6365)."""
    rng = np.random.default_rng(5)
    out = {}
    # pin the fp32 distance formula against torch.norm, incl. points within a few ulp of the cutoff
    v = rng.normal(size=(200000, 3)).astype(np.float32)
    v *= (np.float32(2.0) / np.linalg.norm(v.astype(np.float64), axis=1)).astype(np.float32)[:, None]
    v[::2] = np.nextafter(v[::2], np.float32(3.0))
    tn = torch.norm(torch.from_numpy(v), dim=1).numpy()
    tn1 = np.array([torch.norm(torch.from_numpy(v[k])).item() for k in range(2000)], np.float32)
    mine = G._dist_f32(v)
    assert np.array_equal(tn, mine), "fp32 distance formula differs from torch.norm(dim=1)"
    assert np.array_equal(tn1, mine[:2000]), "fp32 distance formula differs from torch.norm() on a 3-vector"
    print("edges: fp32 distance formula == torch.norm on 200000 near-cutoff vectors; fraction inside:", float((mine < 2.0).mean()))

    def notebook(pos, rc):
        n, e, mn, mi = pos.shape[0], [], 100, None
        p = torch.from_numpy(pos)
        for i in range(n):
            for j in range(n):
                if i != j:
                    d = torch.norm(p[i] - p[j])
                    if d < mn and d > rc:
                        mn, mi = d, [i, j]
                    if d < rc:
                        e.append([i, j])
        if not e:
            e.append(mi)
        return np.array(e, np.int64).T

    clouds = {"dense": (rng.normal(size=(40, 3)) * 2.0).astype(np.float32),
              "sparse": (rng.normal(size=(12, 3)) * 30.0).astype(np.float32)}
    # a lattice whose neighbour distance is exactly / one ulp around the cutoff
    lat = np.stack(np.meshgrid(*[np.arange(3)] * 3, indexing="ij"), -1).reshape(-1, 3).astype(np.float32) * np.float32(2.0)
    lat[1] = np.nextafter(lat[1], np.float32(-9)); lat[5] = np.nextafter(lat[5], np.float32(9))
    clouds["lattice"] = lat
    for k, pos in clouds.items():
        e_ref = notebook(pos, 2.0)
        e_me = G.cutoff_edges(pos, 2.0, None)
        assert np.array_equal(e_ref, e_me), k
        out[f"open_{k}_pos"], out[f"open_{k}_edges"] = pos, e_ref
    # FC order (parts/train_per_iretation.py:308-313)
    for n in (1, 2, 5, 24):
        e = [[i, j] for i in range(n) for j in range(n) if i != j]
        e = np.array(e, np.int64).reshape(-1, 2).T
        assert np.array_equal(e, G.fc_edges(n)), n
        out[f"fc_{n}"] = e
    # periodic builder: defined by graphs_ref (new feature) and cross-checked against 27-image replication
    b = G.make_batch(2, 48, seed=21, edges="cutoff_periodic", r_c=5.0)
    for g in range(2):
        pos = b["pos"][g * 48:(g + 1) * 48]
        box = b["boxes"][g]
        e = G.cutoff_edges(pos, 5.0, box)
        sh = np.array([[a, c, d] for a in (-1, 0, 1) for c in (-1, 0, 1) for d in (-1, 0, 1)], np.float64) * box.astype(np.float64)
        dd = pos[:, None, None, :].astype(np.float64) - pos[None, :, None, :].astype(np.float64) - sh[None, None]
        dmin = np.sqrt((dd ** 2).sum(-1)).min(-1)
        m = dmin < 5.0 - 1e-4
        m2 = dmin < 5.0 + 1e-4
        mine = np.zeros((48, 48), bool); mine[e[0], e[1]] = True
        np.fill_diagonal(m, False); np.fill_diagonal(m2, False)
        assert (mine | ~m).all() and (~mine | m2).all(), "periodic builder disagrees with 27-image replication"
    out["per_pos"], out["per_boxes"], out["per_edges"], out["per_graph_ptr"] = b["pos"], b["boxes"], b["edge_index"], b["graph_ptr"]
    np.savez_compressed(os.path.join(GOLD, "edges.npz"), **out)
    print("edges: open cutoff == notebook loop on", list(clouds), "; FC order ok; periodic ok, E =", b["edge_index"].shape[1])


def golden_loops(egnn_mod, dxh):
    """Loss with injected t/noise and a full T-step trajectory, driven by the UNMODIFIED reference modules."""
    sizes = egnn_ref.net_sizes(4, 64)
    torch.manual_seed(2024)
    ref = egnn_mod.EquivariantGNN(**sizes)
    sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    T = 1000
    sch = D.Schedule(1e-5, 2.0, T)
    dp = dxh.E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=T)
    b = G.make_batch(4, 24, seed=31, edges="fc")
    pos, x1h = torch.from_numpy(b["pos"]), torch.from_numpy(b["x"])
    ei, batch = torch.from_numpy(b["edge_index"]), torch.from_numpy(b["batch"])
    gen = torch.Generator().manual_seed(9)
    t_g = torch.tensor([1, 250, 777, 1000])
    n_pos, n_h = torch.randn(96, 3, generator=gen), torch.randn(96, 2, generator=gen)
    ref.zero_grad()
    loss, aux = LP.train_loss(lambda e, h, x: ref(e, h, x), sch, pos, x1h, ei, batch, t_g, n_pos, n_h)
    loss.backward()
    # the same batch through the reference's own diffuse_zero_to_t with matched RNG: one graph as a spot check
    sel = batch == 2
    torch.manual_seed(1)
    raw = torch.zeros(24, 3).normal_(mean=0, std=1)
    torch.manual_seed(1)
    zt, nz = dp.diffuse_zero_to_t(pos[sel], 777, mode="pos")
    zt2, nz2 = D.diffuse_zero_to_t(sch, pos[sel], 777, raw.clone(), "pos")
    assert torch.equal(zt, zt2) and torch.equal(nz, nz2)
    out = dict(pos=b["pos"], x=b["x"], edge_index=b["edge_index"], batch=b["batch"], t=t_g.numpy(), noise_pos=n_pos.numpy(),
               noise_h=n_h.numpy(), loss=np.float64(loss.item()), pos_t=aux["pos_t"].numpy(), h_in=aux["h_in"].detach().numpy(),
               eps_pos=aux["eps_pos"].numpy(), h_out=aux["h_out"].detach().numpy(), x_out=aux["x_out"].detach().numpy())
    for k, v in sd.items():
        out["w:" + k] = v.numpy()
    for k, p in ref.named_parameters():
        out["g:" + k] = p.grad.numpy()
    np.savez_compressed(os.path.join(GOLD, "loss_cfg1.npz"), **out)
    print("loss_cfg1: loss =", loss.item())

    # trajectory: one 24-atom structure, T=1000.  Untrained weights diverge under the production schedule
    # (s=1e-5 => z grows 1e5-fold), so the trajectory fixture uses tamed weights and noise_precision s=0.05;
    # the maths per step is identical.  Noise comes from a seeded CPU generator (regenerated in the test).
    n = 24
    sd_t = egnn_ref.tame_state_dict(sd, sizes["L"])
    ref.load_state_dict(sd_t)
    sch_t = D.Schedule(0.05, 2.0, T)
    noises = LP.make_noises(n, T, 2, torch.Generator().manual_seed(123))
    eifc = torch.from_numpy(G.fc_edges(n))
    rec = {}
    with torch.no_grad():
        pos_f, h_f, atoms, finite = LP.sample_structure(lambda e, h, x: ref(e, h, x), sch_t, eifc, n, noises,
                                                        record=lambda t, zx, zh: rec.__setitem__(t, (zx.clone(), zh.clone())))
        # sensitivity of the trajectory to round-off: same loop in float64 through the restatement
        sd64 = {k: v.double() for k, v in sd_t.items()}
        n64 = {k: v.double() for k, v in noises.items()}
        sch64 = D.Schedule(0.05, 2.0, T); sch64.alpha = sch_t.alpha.double(); sch64.sigma = sch_t.sigma.double()
        pos64, h64, _, _ = LP.sample_structure(lambda e, h, x: egnn_ref.egnn_forward(sd64, sizes["L"], e, h.double(), x),
                                               sch64, eifc, n, n64)
    dev = relerr(pos_f.double(), pos64)
    keep = [1000, 999, 900, 500, 100, 1]
    out = dict(seed=np.int64(123), n=np.int64(n), T=np.int64(T), s=np.float64(0.05), pos=pos_f.numpy(), h=h_f.numpy(),
               atoms=atoms.numpy(), finite=np.bool_(finite), keep=np.array(keep), init_x=noises["init_x"].numpy(),
               step_x_first=noises["step_x"][0].numpy(), fp32_vs_fp64_relerr=np.float64(dev))
    for t in keep:
        out[f"zx_t{t}"], out[f"zh_t{t}"] = rec[t][0].numpy(), rec[t][1].numpy()
    np.savez_compressed(os.path.join(GOLD, "traj_cfg1.npz"), **out)
    print("traj_cfg1: finite =", finite, "|pos|max =", float(pos_f.abs().max()), "fp32-vs-fp64 relerr =", dev)


def golden_conditional(egnn_mod):
    """Conditional model (parameters.yaml: conditional, to_compress_spectrum, give_exO): h = [h_t | compressor(spectrum) |
    exO | t/T] (parts/train_per_iretation.py:136-145), H = 2 + 32 + 1 + 1 = 36, driven by the UNMODIFIED reference EGNN
    and DataPreprocessor.SpectrumCompressor; widths reduced so the fixture stays small."""
    import importlib
    dpp = importlib.import_module("DataPreprocessor")
    A, C, S = 2, 32, 200
    H = A + C + 1 + 1
    sizes = dict(L=2, m_input=2 * H + 1, m_hidden=96, m_output=64, x_input=2 * H + 1, x_hidden=80, x_output=1,
                 h_input=H + 64, h_hidden=112, h_output=H)
    torch.manual_seed(2024)
    ref = egnn_mod.EquivariantGNN(**sizes)
    comp = dpp.SpectrumCompressor(S, [64, 48], C)
    sd = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    csd = {k: v.detach().clone() for k, v in comp.state_dict().items()}
    T = 1000
    sch = D.Schedule(1e-5, 2.0, T)
    b = G.make_batch(3, 20, seed=41, edges="fc")
    pos, x1h = torch.from_numpy(b["pos"]), torch.from_numpy(b["x"])
    ei, batch = torch.from_numpy(b["edge_index"]), torch.from_numpy(b["batch"])
    N = pos.shape[0]
    gen = torch.Generator().manual_seed(17)
    spectrum = torch.rand(N, S, generator=gen)
    exO = (torch.arange(N) % 20 == 0).float().view(N, 1)  # the excited O is atom 0 of every structure
    t_g = torch.tensor([3, 512, 1000])
    n_pos, n_h = torch.randn(N, 3, generator=gen), torch.randn(N, A, generator=gen)
    cond = torch.cat((comp(spectrum), exO), dim=1)
    loss, aux = LP.train_loss(lambda e, h, x: ref(e, h, x), sch, pos, x1h, ei, batch, t_g, n_pos, n_h, A=A, cond=cond)
    loss.backward()
    out = dict(pos=b["pos"], x=b["x"], edge_index=b["edge_index"], batch=b["batch"], t=t_g.numpy(), noise_pos=n_pos.numpy(),
               noise_h=n_h.numpy(), spectrum=spectrum.numpy(), exO=exO.numpy(), loss=np.float64(loss.item()),
               h_in=aux["h_in"].detach().numpy(), h_out=aux["h_out"].detach().numpy(), x_out=aux["x_out"].detach().numpy(),
               size_keys=np.array(list(sizes.keys())), sizes=np.array(list(sizes.values())))
    for k, v in sd.items():
        out["w:" + k] = v.numpy()
    for k, p_ in ref.named_parameters():
        out["g:" + k] = p_.grad.numpy()
    for k, v in csd.items():
        out["cw:" + k] = v.numpy()
    for k, p_ in comp.named_parameters():
        out["cg:" + k] = p_.grad.numpy()
    np.savez_compressed(os.path.join(GOLD, "loss_cond.npz"), **out)
    print("loss_cond: H =", H, "loss =", loss.item())


YAML_SIZES = dict(L=2, m_input=73, m_hidden=1024, m_output=256, x_input=73, x_hidden=1024, x_output=1, h_input=36 + 256,
                  h_hidden=1024, h_output=36)
YAML_SEED = 2024
YAML_SUB = 8  # gradients of matrices with more than 50 000 entries are stored as g[::8, ::8] (and their full norm)


def yaml_modules(egnn_cls, compressor_cls):
    """The seeded network of the 'loss_yaml' golden: parameters.yaml widths (main.py:102-121), two layers, default init.
    The fixture stores a checksum instead of 3.6 M weights; the test rebuilds them with the same seed and class order."""
    torch.manual_seed(YAML_SEED)
    net = egnn_cls(**YAML_SIZES)
    comp = compressor_cls(200, [150, 100, 50], 32)
    return net, comp


def weight_checksum(sd):
    return np.array([float(v.double().sum()) for v in sd.values()] + [float(v.double().square().sum()) for v in sd.values()])


def golden_yaml(egnn_mod):
    """The reference's own configuration (parameters.yaml: conditional, compressed spectrum 200 -> 150 -> 100 -> 50 -> 32, exO,
    m/x/h hidden 1024, m_size 256; H = 36, K0 = 73) at L = 2 on two fully connected 14-atom structures: loss, outputs and
    gradients of the eps-prediction training loss from the UNMODIFIED reference EGNN and SpectrumCompressor."""
    import importlib
    dpp = importlib.import_module("DataPreprocessor")
    A, S = 2, 200
    ref, comp = yaml_modules(egnn_mod.EquivariantGNN, dpp.SpectrumCompressor)
    sch = D.Schedule(1e-5, 2.0, 1000)
    b = G.make_batch(2, 14, seed=43, edges="fc")
    pos, x1h = torch.from_numpy(b["pos"]), torch.from_numpy(b["x"])
    ei, batch = torch.from_numpy(b["edge_index"]), torch.from_numpy(b["batch"])
    N = pos.shape[0]
    gen = torch.Generator().manual_seed(19)
    spectrum = torch.rand(N, S, generator=gen)
    exO = (torch.arange(N) % 14 == 0).float().view(N, 1)
    t_g = torch.tensor([40, 870])
    n_pos, n_h = torch.randn(N, 3, generator=gen), torch.randn(N, A, generator=gen)
    cond = torch.cat((comp(spectrum), exO), dim=1)
    loss, aux = LP.train_loss(lambda e, h, x: ref(e, h, x), sch, pos, x1h, ei, batch, t_g, n_pos, n_h, A=A, cond=cond)
    loss.backward()
    out = dict(pos=b["pos"], x=b["x"], edge_index=b["edge_index"], batch=b["batch"], t=t_g.numpy(), noise_pos=n_pos.numpy(),
               noise_h=n_h.numpy(), spectrum=spectrum.numpy(), exO=exO.numpy(), loss=np.float64(loss.item()),
               h_in=aux["h_in"].detach().numpy(), h_out=aux["h_out"].detach().numpy(), x_out=aux["x_out"].detach().numpy(),
               wsum=weight_checksum(ref.state_dict()), cwsum=weight_checksum(comp.state_dict()))
    for mod, tag in ((ref, "g:"), (comp, "cg:")):
        for k, p_ in mod.named_parameters():
            g = p_.grad
            out[tag + "norm:" + k] = np.float64(g.double().norm().item())
            out[tag + k] = (g[::YAML_SUB, ::YAML_SUB] if g.numel() > 50000 else g).numpy()
    np.savez_compressed(os.path.join(GOLD, "loss_yaml.npz"), **out)
    print("loss_yaml: loss =", loss.item(), "|h_out| max", float(aux["h_out"].abs().max()))


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(4)
    egnn_mod, dxh, dnew = load_reference()
    golden_diffusion(dxh, dnew)
    golden_edges()
    golden_egnn(egnn_mod)
    golden_loops(egnn_mod, dxh)
    golden_conditional(egnn_mod)
    golden_yaml(egnn_mod)
    print("golden vectors written to", GOLD)


if __name__ == "__main__":
    main()
