"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C ABI, against the golden
vectors frozen from the unmodified reference and against the CPU oracle on the same seeded inputs.

Bars: bit-exact for edge lists and index structures; fp32 mode max|err|/max|ref| <= 1e-4 per layer (BASELINE north_star).
"""
import os

import numpy as np
import pytest
import torch

from oracle import diffusion_ref as D
from oracle import egnn_ref, graphs_ref as G, loops_ref as LP

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 1e-4


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def state_dict_of(z):
    return {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w:")}


def sizes_of(z):
    return {str(k): int(v) for k, v in zip(z["size_keys"], z["sizes"])}


def relerr(a, b):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def make_net(z):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    net = EquivariantGNN(**sizes_of(z))
    net.load_state_dict(state_dict_of(z))
    net.precision = "fp32"  # these tests pin the CUDA-core fp32 mode; the tcgen05 modes are pinned in test_gpu_tc.py
    return net.to(DEV)


# ---------------------------------------------------------------- edge builders: bit exact
def test_fc_edges_bit_exact(lib, golden_dir):
    from diffusion_model_b200.graph import build_edges_fc
    z = load(golden_dir, "edges")
    for n in (2, 5, 24):
        assert np.array_equal(build_edges_fc([n]).cpu().numpy(), z[f"fc_{n}"].reshape(2, -1))
    assert build_edges_fc([1]).shape == (2, 0)
    sizes = [3, 1, 24, 7, 192]
    assert np.array_equal(build_edges_fc(sizes).cpu().numpy(), G.fc_edges_batch(sizes))


def test_cutoff_edges_bit_exact(lib, golden_dir):
    from diffusion_model_b200.graph import build_edges_cutoff
    z = load(golden_dir, "edges")
    for k in ("dense", "lattice"):
        e = build_edges_cutoff(torch.from_numpy(z[f"open_{k}_pos"]).to(DEV), 2.0)
        assert np.array_equal(e.cpu().numpy(), z[f"open_{k}_edges"]), k
    e = build_edges_cutoff(torch.from_numpy(z["per_pos"]).to(DEV), 5.0,
                           node_ptr=torch.from_numpy(z["per_graph_ptr"]), box=torch.from_numpy(z["per_boxes"]))
    assert np.array_equal(e.cpu().numpy(), z["per_edges"])
    # near-cutoff stress: pairs within a few ulp of r_c, oracle as the checker
    rng = np.random.default_rng(3)
    v = rng.normal(size=(300, 3)).astype(np.float32)
    v *= (np.float32(2.0) / np.linalg.norm(v.astype(np.float64), axis=1)).astype(np.float32)[:, None]
    v[::2] = np.nextafter(v[::2], np.float32(3.0))
    pos = np.concatenate([np.zeros((1, 3), np.float32), v])
    e = build_edges_cutoff(torch.from_numpy(pos).to(DEV), 2.0)
    assert np.array_equal(e.cpu().numpy(), G.cutoff_edges(pos, 2.0, None, fallback=False))
    # full-size config-2 shaped batch (64 x 192 atoms, periodic 5 A): against the oracle
    b = G.make_batch(8, 192, seed=5, edges="cutoff_periodic")
    e = build_edges_cutoff(torch.from_numpy(b["pos"]).to(DEV), 5.0, node_ptr=torch.from_numpy(b["graph_ptr"]),
                           box=torch.from_numpy(b["boxes"]))
    assert np.array_equal(e.cpu().numpy(), b["edge_index"])


def test_graph_index_structures(lib):
    from diffusion_model_b200.graph import GraphIndex
    from diffusion_model_b200 import EgnnError
    b = G.make_batch(3, 40, seed=9, edges="cutoff_open", r_c=4.0)
    ei = torch.from_numpy(b["edge_index"]).to(DEV)
    N, E = b["pos"].shape[0], ei.shape[1]
    gi = GraphIndex(ei, N, batch=torch.from_numpy(b["batch"]).to(DEV))
    recv, send = b["edge_index"]
    assert np.array_equal(gi.recv[:E].cpu().numpy(), recv) and np.array_equal(gi.send[:E].cpu().numpy(), send)
    row_ptr = np.searchsorted(recv, np.arange(N + 1))
    assert np.array_equal(gi.row_ptr.cpu().numpy(), row_ptr)
    order = np.argsort(send, kind="stable")
    col_ptr = np.searchsorted(send[order], np.arange(N + 1))
    assert np.array_equal(gi.col_ptr.cpu().numpy(), col_ptr)
    assert np.array_equal(gi.col_edge[:E].cpu().numpy(), order)
    assert np.array_equal(gi.node_ptr.cpu().numpy(), b["graph_ptr"])
    # any edge order is legal for the reference's scatter-add (PyG propagate, :68,:70): an unsorted list is stable-sorted
    # by receiver on the device, the derived structures are those of the sorted list, `perm` maps kernel edges to input edges
    rev = ei.flip(1).contiguous()
    gu = GraphIndex(rev, N)
    perm = torch.sort(rev[0], stable=True).indices
    assert torch.equal(gu.perm, perm)
    srt = rev[:, perm].cpu().numpy()
    assert np.array_equal(gu.recv[:E].cpu().numpy(), srt[0]) and np.array_equal(gu.send[:E].cpu().numpy(), srt[1])
    assert np.array_equal(gu.row_ptr.cpu().numpy(), row_ptr)
    assert np.array_equal(gu.col_ptr.cpu().numpy(), col_ptr)
    with pytest.raises(EgnnError):
        GraphIndex(ei + 1000, N)   # ids outside [0, N) still raise
    gi0 = GraphIndex(torch.zeros(2, 0, dtype=torch.int64, device=DEV), 5)  # empty edge list
    assert gi0.row_ptr.tolist() == [0] * 6


# ---------------------------------------------------------------- EGNN forward, fp32 mode
@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg4", "ragged"])
def test_egnn_forward_matches_reference_layers(lib, golden_dir, name):
    z = load(golden_dir, "egnn_" + name)
    net = make_net(z)
    L = sizes_of(z)["L"]
    h = torch.from_numpy(z["h"]).to(DEV)
    x = torch.from_numpy(z["x"]).to(DEV)
    ei = torch.from_numpy(z["edge_index"]).to(DEV)
    batch = torch.from_numpy(z["batch"]).to(DEV)
    with torch.no_grad():
        hh, xx = h, x
        for l in range(L):  # layer by layer through the EGCL drop-in, fed with the reference's previous output
            hh, xx = net.egcl_list[l](ei, torch.from_numpy(z[f"h_l{l - 1}"]).to(DEV) if l else h,
                                      torch.from_numpy(z[f"x_l{l - 1}"]).to(DEV) if l else x)
            assert relerr(hh, z[f"h_l{l}"]) <= TOL, (name, l, relerr(hh, z[f"h_l{l}"]))
            assert relerr(xx, z[f"x_l{l}"]) <= TOL, (name, l, relerr(xx, z[f"x_l{l}"]))
        ho, xo = net(ei, h, x)
        assert relerr(ho, z[f"h_l{L - 1}"]) <= TOL and relerr(xo, z[f"x_l{L - 1}"]) <= TOL
        h2, x2 = net(h, x, ei)  # north_star argument order
        assert torch.equal(h2, ho) and torch.equal(x2, xo)
        hg, xg = net(ei, h, x, batch=batch, norm_scope="graph")
        assert relerr(hg, z["h_graph"]) <= TOL and relerr(xg, z["x_graph"]) <= TOL
        # bit-identical reruns (no atomics, fixed reduction order)
        h3, x3 = net(ei, h, x)
        assert torch.equal(h3, ho) and torch.equal(x3, xo)


def test_egnn_forward_edge_cases(lib):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    torch.manual_seed(0)
    sz = egnn_ref.net_sizes(2, 32)
    net = EquivariantGNN(**sz).to(DEV)
    net.precision = "fp32"
    sd = {k: v.detach().cpu() for k, v in net.state_dict().items()}
    gen = torch.Generator().manual_seed(1)
    # (a) isolated nodes + a 300-edge receiver crossing tile boundaries; (b) no edges at all
    n = 320
    recv = np.concatenate([np.full(300, 2), np.full(5, 7), np.full(200, 9)]).astype(np.int64)
    send = np.concatenate([np.arange(10, 310), np.arange(0, 5), np.arange(100, 300)]).astype(np.int64)
    for ei in (torch.from_numpy(np.stack([recv, send])), torch.zeros(2, 0, dtype=torch.int64)):
        h, x = torch.randn(n, 3, generator=gen), torch.randn(n, 3, generator=gen)
        with torch.no_grad():
            ho, xo = net(ei.to(DEV), h.to(DEV), x.to(DEV))
            hr, xr = egnn_ref.egnn_forward(sd, 2, ei, h, x)
        assert relerr(ho, hr) <= TOL and relerr(xo, xr) <= TOL


# ---------------------------------------------------------------- diffusion kernels
def test_remove_mean(lib, golden_dir):
    from diffusion_model_b200.diffusion_x_h import remove_mean
    z = load(golden_dir, "diffusion_kat")
    x = torch.from_numpy(z["rm_in"]).to(DEV)
    out = remove_mean(x, torch.from_numpy(z["rm_batch"]).to(DEV))
    assert out.data_ptr() == x.data_ptr()  # in place, like the reference
    assert np.array_equal(out.cpu().numpy(), z["rm_out"])
    assert np.array_equal(remove_mean(torch.from_numpy(z["rm_in"]).to(DEV)).cpu().numpy(), z["rm_out_global"])
    gen = torch.Generator().manual_seed(4)
    big = torch.randn(1000, 3, generator=gen)
    b = torch.repeat_interleave(torch.arange(5), 200)
    ref = D.remove_mean(big.clone(), b)
    got = remove_mean(big.clone().to(DEV), b.to(DEV))
    assert relerr(got, ref) < 1e-6


def test_schedule_and_step_scalars(lib, golden_dir):
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    z = load(golden_dir, "diffusion_kat")
    dp = E3DiffusionProcess(s=float(z["s"]), power=float(z["power"]), num_diffusion_timestep=int(z["T"]))
    assert np.array_equal(dp.alpha_schedule.numpy(), z["alpha"]) and np.array_equal(dp.sigma_schedule.numpy(), z["sigma"])
    _, _, coef = dp.device_tables(DEV)
    for t in z["ts"].tolist():
        assert np.array_equal(coef[t, :4].cpu().numpy(), z[f"t{t}_scalars"])
    assert np.array_equal(coef[1:, 4].cpu().numpy(), z["time_col"][:-1])


def test_sampler_step_kernel(lib, golden_dir):
    from diffusion_model_b200._lib import call, ptr, stream_ptr
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    z = load(golden_dir, "diffusion_kat")
    T = int(z["T"])
    sch = D.Schedule(float(z["s"]), float(z["power"]), T)
    dp = E3DiffusionProcess(s=float(z["s"]), power=float(z["power"]), num_diffusion_timestep=T)
    _, _, coef = dp.device_tables(DEV)
    gen = torch.Generator().manual_seed(8)
    S, n, A, H = 3, 24, 2, 3
    N = S * n
    node_ptr = (torch.arange(S + 1) * n).to(torch.int32).to(DEV)
    batch = torch.repeat_interleave(torch.arange(S), n)
    for t in (1000, 500, 1):
        z_x, x_out = torch.randn(N, 3, generator=gen), torch.randn(N, 3, generator=gen)
        h_in, h_out = torch.randn(N, H, generator=gen), torch.randn(N, H, generator=gen)
        nx, nh = torch.randn(N, 3, generator=gen), torch.randn(N, A, generator=gen)
        # oracle, structure by structure
        ex, eh_ = [], []
        for s_ in range(S):
            sl = slice(s_ * n, (s_ + 1) * n)
            eps = D.remove_mean(x_out[sl] - z_x[sl])
            ex.append(D.reverse_diffuse_one_step(sch, z_x[sl], eps, t, nx[sl].clone(), "pos"))
            eh_.append(D.reverse_diffuse_one_step(sch, h_in[sl, :A], h_out[sl, :A], t, nh[sl].clone(), "h"))
        zx_d, hin_d = z_x.to(DEV), h_in.to(DEV)
        xo_d, ho_d, nx_d, nh_d = x_out.to(DEV), h_out.to(DEV), nx.to(DEV), nh.to(DEV)  # keep alive across the launch
        finite = torch.ones(S, dtype=torch.int32, device=DEV)
        t_dev = torch.tensor([t], dtype=torch.int32, device=DEV)
        call("egnn_sampler_step", ptr(zx_d), ptr(hin_d), ptr(xo_d), ptr(ho_d), ptr(nx_d), ptr(nh_d), ptr(coef),
             ptr(t_dev), t, 0, 0, 0, 1.0, ptr(node_ptr), S, H, A, 0, None, ptr(finite), stream_ptr())
        assert torch.equal(hin_d[:, :A].cpu(), torch.cat(eh_))  # h mode: bit exact (no reduction involved)
        assert relerr(zx_d, torch.cat(ex)) < 2e-6
        assert torch.equal(hin_d[:, H - 1].cpu(), torch.full((N,), (t - 1) / T, dtype=torch.float32))
        assert finite.tolist() == [1] * S
    # non-finite input of structure 1 clears only its flag
    zx_d = torch.randn(N, 3, generator=gen).to(DEV)
    zx_d[n + 3, 0] = float("nan")
    finite = torch.ones(S, dtype=torch.int32, device=DEV)
    call("egnn_sampler_step", ptr(zx_d), ptr(hin_d), ptr(xo_d), ptr(ho_d), ptr(nx_d), ptr(nh_d), ptr(coef),
         ptr(t_dev), 1, 0, 0, 0, 1.0, ptr(node_ptr), S, H, A, 0, None, ptr(finite), stream_ptr())
    assert finite.tolist() == [1, 0, 1]


def test_diffuse_as_batch_against_oracle(lib, golden_dir):
    from diffusion_model_b200.data import Batch, Data
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.train_per_iretation import diffuse_as_batch
    z = load(golden_dir, "loss_cfg1")
    dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=1000)
    sch = D.Schedule(1e-5, 2.0, 1000)
    pos, x1h, batch = torch.from_numpy(z["pos"]), torch.from_numpy(z["x"]), torch.from_numpy(z["batch"])
    bd = Data(x=x1h.to(DEV), pos=pos.to(DEV), edge_index=torch.from_numpy(z["edge_index"]).to(DEV))
    bd.batch = batch.to(DEV)
    t = z["t"].tolist()
    out = diffuse_as_batch(bd, bd.batch, dict(num_diffusion_timestep=1000, spectrum_size=0, onehot_scaling_factor=1.0),
                           dp, t=t, noise_pos=torch.from_numpy(z["noise_pos"]).to(DEV),
                           noise_h=torch.from_numpy(z["noise_h"]).to(DEV))
    assert relerr(out.pos_at_t, z["pos_t"]) < 1e-6
    assert relerr(out.y_for_noise_pos, z["eps_pos"]) < 1e-6
    assert torch.equal(out.h_at_t.cpu(), torch.from_numpy(z["h_in"][:, :2]))  # bit exact: no reduction
    assert out.each_time_list.shape == (96, 1) and out.each_time_list[:, 0].cpu().tolist() == sum([[q] * 24 for q in t], [])


def test_trajectory_matches_reference(lib, golden_dir):
    """T = 1000 ancestral steps of one 24-atom structure with injected noise, against the trajectory the unmodified
    reference modules produced (oracle/make_golden.py: tamed weights, s = 0.05)."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.sampler import sample_structures
    zt = load(golden_dir, "traj_cfg1")
    zl = load(golden_dir, "loss_cfg1")
    sd = egnn_ref.tame_state_dict(state_dict_of(zl), 4)
    net = EquivariantGNN(**egnn_ref.net_sizes(4, 64))
    net.load_state_dict(sd)
    net = net.to(DEV)
    net.precision = "fp32"
    n, T = int(zt["n"]), int(zt["T"])
    dp = E3DiffusionProcess(s=float(zt["s"]), power=2.0, num_diffusion_timestep=T)
    noises = LP.make_noises(n, T, 2, torch.Generator().manual_seed(int(zt["seed"])))
    assert np.array_equal(noises["init_x"].numpy(), zt["init_x"])
    keep = [int(k) for k in zt["keep"]]
    for graphs in (False, True):
        out = sample_structures(net, dp, n, 1, noises=noises, record_steps=() if graphs else keep,
                                use_cuda_graph=graphs)
        tol = max(50 * float(zt["fp32_vs_fp64_relerr"]), 5e-4)  # the trajectory's own fp32 sensitivity sets the bar
        assert bool(out["finite"][0]) == bool(zt["finite"])
        assert relerr(out["pos"][0], zt["pos"]) < tol, relerr(out["pos"][0], zt["pos"])
        assert relerr(out["h"][0], zt["h"]) < tol
        assert np.array_equal(out["atoms"][0].cpu().numpy(), zt["atoms"])
        for t, (zx, zh) in out["record"].items():
            assert relerr(zx, zt[f"zx_t{t}"]) < tol and relerr(zh, zt[f"zh_t{t}"]) < tol, t


def test_deferred_edge_list_validation(lib):
    """training loops validate the edge list without a host sync per minibatch; the error surfaces at check_pending()"""
    from diffusion_model_b200 import EgnnError
    from diffusion_model_b200.graph import GraphIndex, check_pending
    good = torch.tensor([[0, 0, 1, 2], [1, 2, 0, 0]], device=DEV)
    bad = torch.tensor([[1, 0, 2, 0], [0, 1, 0, 2]], device=DEV)   # receivers not sorted
    oob = torch.tensor([[0, 0, 1, 2], [1, 7, 0, 0]], device=DEV)   # sender id out of range
    GraphIndex(good, 3, validate="deferred")
    check_pending()
    for ei in (bad, oob):
        gi = GraphIndex(ei, 3, validate="deferred")
        with pytest.raises(EgnnError):
            check_pending()
        assert gi._pending is None
    check_pending()  # nothing left behind: a reported list never resurfaces in an unrelated caller
    from diffusion_model_b200.graph import _PENDING
    assert not _PENDING
    # the owner of a malformed deferred list keeps getting the error; nobody else does
    gi = GraphIndex(bad, 3, validate="deferred")
    with pytest.raises(EgnnError):
        gi.check()
    with pytest.raises(EgnnError):
        gi.check()
    check_pending()
    assert GraphIndex(bad, 3).perm is not None   # immediate validation sorts an unsorted list instead of raising
    with pytest.raises(EgnnError):
        GraphIndex(oob, 3)


def test_xonly_variant_against_oracle(lib, golden_dir):
    """E3diffusion_new drop-in (train.py / test.py path): posterior mean in x_hat form, COM-free
    noise, and one train.py-style step (loss + gradients) against the oracle restatement"""
    from diffusion_model_b200.E3diffusion_new import E3DiffusionProcess, remove_mean
    z = load(golden_dir, "diffusion_kat")
    T = int(z["T"])
    dp = E3DiffusionProcess(s=float(z["s"]), power=float(z["power"]), num_diffusion_timestep=T)
    sch = D.Schedule(float(z["s"]), float(z["power"]), T)
    zz, eps = torch.from_numpy(z["z"]), torch.from_numpy(z["eps"])
    for t in (1000, 500, 1):
        mu = dp.calculate_mu(zz.to(DEV), eps.to(DEV), t)
        assert relerr(mu, z[f"t{t}_mu_xonly"]) <= 1e-6
        torch.manual_seed(t)
        nxt = dp.reverse_diffuse_one_step(mu, t)
        step = (nxt - mu).cpu()
        assert float(step.mean(dim=0).abs().max()) <= 1e-5 * max(1.0, float(step.abs().max()))  # COM-free noise
    # one training step of train.py:238-288 on a 24-atom fully connected graph
    zl = load(golden_dir, "loss_cfg1")
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    sd = state_dict_of(zl)
    net = EquivariantGNN(**egnn_ref.net_sizes(4, 64))
    net.load_state_dict(sd)
    net = net.to(DEV)
    net.precision = "fp32"
    n, t = 24, 400
    gen = torch.Generator().manual_seed(5)
    pos = D.remove_mean(torch.randn(n, 3, generator=gen) * 2)
    onehot = torch.nn.functional.one_hot(torch.randint(0, 2, (n,), generator=gen), 2).float()
    ei = torch.as_tensor(G.fc_edges(n))
    torch.manual_seed(7)
    pos_t, noise = dp.diffuse_zero_to_t(pos.to(DEV), t)
    assert relerr(pos_t, sch.alpha[t] * pos + sch.sigma[t] * noise.cpu()) <= 1e-6
    h = torch.cat((onehot.to(DEV), torch.full((n, 1), t / T, device=DEV)), dim=1)
    _, x = net(ei.to(DEV), h, pos_t)
    loss = ((remove_mean(x - pos_t) - noise) ** 2).sum()
    loss.backward()
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ref = LP.train_loss_xonly(lambda e, hh, xx: egnn_ref.egnn_forward(sdg, 4, e, hh, xx), sch, pos, onehot, ei, t,
                              noise.cpu())
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 2e-5 * abs(ref.item())
    for k, p in net.named_parameters():
        if sdg[k].grad is not None:
            scale = max(float(sdg[k].grad.abs().max()), 1e-3 * max(float(v.grad.abs().max()) for v in sdg.values()
                                                                   if v.grad is not None))
            assert float((p.grad.cpu() - sdg[k].grad).abs().max()) <= 1e-3 * scale, k


def test_cutoff_fallback_closest_pair(lib):
    """graphs with no pair inside the cutoff get the reference loop's closest ordered pair (something.ipynb:6346-6358):
    first strict minimum in row-major order, ties and pairs exactly at the cutoff included; batched with normal graphs"""
    from diffusion_model_b200.graph import build_edges_cutoff
    rng = np.random.default_rng(11)
    graphs = [
        rng.normal(size=(12, 3)).astype(np.float32) * 1.0,                      # normal graph with edges
        (rng.normal(size=(9, 3)) * 40).astype(np.float32),                      # sparse: no pair inside 2.0
        np.array([[0, 0, 0], [3, 0, 0], [0, 3, 0], [-3, 0, 0]], np.float32),    # tie: several pairs at distance 3
        np.array([[0, 0, 0], [2, 0, 0], [7, 0, 0]], np.float32),                # a pair exactly AT the cutoff is skipped
        np.array([[0, 0, 0], [500, 0, 0]], np.float32),                         # nothing below 100: stays empty
        np.array([[1, 2, 3]], np.float32),                                      # single atom
        rng.normal(size=(6, 3)).astype(np.float32) * 0.5,
    ]
    ptr_ = np.concatenate([[0], np.cumsum([len(g) for g in graphs])]).astype(np.int32)
    want = np.concatenate([G.cutoff_edges(g, 2.0, None, offset=int(o)) for g, o in zip(graphs, ptr_[:-1])], axis=1)
    pos = torch.from_numpy(np.concatenate(graphs)).to(DEV)
    got = build_edges_cutoff(pos, 2.0, node_ptr=torch.from_numpy(ptr_))
    assert np.array_equal(got.cpu().numpy(), want)
    assert want.shape[1] > 0 and np.array_equal(want[:, (want[0] >= 12) & (want[0] < 21)].shape, (2, 1))
    plain = build_edges_cutoff(pos, 2.0, node_ptr=torch.from_numpy(ptr_), fallback=False)
    nofb = np.concatenate([G.cutoff_edges(g, 2.0, None, offset=int(o), fallback=False) for g, o in zip(graphs, ptr_[:-1])], axis=1)
    assert np.array_equal(plain.cpu().numpy(), nofb)
    # periodic: minimum-image distances decide the fallback pair too
    box = np.array([[30.0, 30.0, 30.0]], np.float32)
    p = np.array([[1, 1, 1], [28, 1, 1], [15, 15, 15]], np.float32)   # atoms 0 and 1 are 3 apart through the boundary
    got = build_edges_cutoff(torch.from_numpy(p).to(DEV), 2.0, node_ptr=torch.tensor([0, 3]), box=torch.from_numpy(box))
    assert np.array_equal(got.cpu().numpy(), G.cutoff_edges(p, 2.0, box[0]))
    assert got.cpu().tolist() == [[0], [1]]
