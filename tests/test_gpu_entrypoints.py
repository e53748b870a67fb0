"""The public entry points a user of the reference calls, each against the CPU oracle under IDENTICAL random draws:

  * train_per_iretation.generate()      (parts/train_per_iretation.py:264-444)
  * train_per_iretation.eval_epoch()    (:185-262) and train_epoch()'s returned per-node loss (:99-183)
  * E3DiffusionProcess.loss() / .sample()  (the BASELINE.json wording, aliases of batch_loss / sample_structures)

The loops draw their own randomness (python `random.choice` for t, the device generator for noise), so each test seeds
both generators, runs the entry point, re-seeds, replays the same draw sequence (same shapes, same order, same device)
and hands those draws to oracle/loops_ref.py.
"""
import random

import numpy as np
import pytest
import torch

from oracle import diffusion_ref as D
from oracle import egnn_ref, graphs_ref as G, loops_ref as LP

pytestmark = pytest.mark.gpu
DEV = "cuda"
A = 2


def relerr(a, b):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _tamed_net(L, d, seed=0):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    torch.manual_seed(seed)
    sizes = egnn_ref.net_sizes(L, d)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), L)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    return net.to(DEV), sd, sizes


def _dataset(num, seed=5):
    from diffusion_model_b200.data import Data
    gen = torch.Generator().manual_seed(seed)
    data = []
    for g in range(num):
        n = 14 + 3 * g
        pos = 4.0 * torch.rand(n, 3, generator=gen)
        x = torch.nn.functional.one_hot(torch.randint(0, 2, (n,), generator=gen), 2).float()
        data.append(Data(x=x, pos=pos, edge_index=torch.as_tensor(G.fc_edges(n)).long(), id=g))
    return data


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-4), ("bf16x3", 1e-3)])
def test_generate_matches_oracle_under_identical_noise(lib, precision, tol):
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.train_per_iretation import generate
    T, S, seed = 40, 3, 1234
    net, sd, sizes = _tamed_net(3, 64)
    net.precision = precision
    # s = 0.05: an untrained net's trajectory grows by 1 / alpha_T = 20; with s = 1e-3 it would be 1000 x and every sample
    # would trip generate()'s `pos > 1000` reject (:434), which redraws for ever -- in the reference's loop as well
    dp = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=T)
    sch = D.Schedule(0.05, 2.0, T)
    test_data = _dataset(2)
    params = dict(atom_type_size=A, onehot_scaling_factor=1.0, conditional=False, give_exO=False)
    torch.manual_seed(seed)
    orig, gen_list = generate({"egnn": net}, test_data, params, dp, gen_num_per_spectrum=S)
    assert len(orig) == len(gen_list) == 2 * S and all(o == -1 for o in orig)  # unconditional: -1 placeholders (:440)
    # replay the device generator: per test graph init_x, init_h, the warm-up fill and the t = T fill of the noise ring
    # (chunk = min(50, T) steps), final_x, final_h -- the draw order of sampler.sample_structures
    torch.manual_seed(seed)
    k = 0
    for data in test_data:
        n = data.x.shape[0]
        N = S * n
        init_x = torch.randn(N, 3, device=DEV)
        init_h = torch.randn(N, A, device=DEV)
        for _ in range(2):
            step_x = torch.empty(T, N, 3, device=DEV).normal_()
            step_h = torch.empty(T, N, A, device=DEV).normal_()
        final_x = torch.randn(N, 3, device=DEV)
        final_h = torch.randn(N, A, device=DEV)
        ei = torch.as_tensor(G.fc_edges(n)).long()
        fwd = lambda e, h, x: egnn_ref.egnn_forward(sd, sizes["L"], e, h, x)
        for q in range(S):
            sl = slice(q * n, (q + 1) * n)
            noises = dict(init_x=init_x[sl].cpu(), init_h=init_h[sl].cpu(), step_x=step_x[:, sl].cpu(),
                          step_h=step_h[:, sl].cpu(), final_x=final_x[sl].cpu(), final_h=final_h[sl].cpu())
            with torch.no_grad():
                pos, hf, atoms, finite = LP.sample_structure(fwd, sch, ei, n, noises)
            traj = gen_list[k]
            assert len(traj) == T // 100 + 1 and all(g is traj[0] for g in traj)  # aliases of ONE Data object (:336, :432)
            g = traj[-1]
            assert finite and relerr(g.pos, pos) <= tol, (precision, k, relerr(g.pos, pos))
            assert relerr(g.h, hf) <= tol
            assert torch.equal(g.x.cpu(), atoms) and torch.equal(g.edge_index.cpu(), ei)
            k += 1


def _replay_epoch(sd, sizes, loader, sch, T, train, lr=1e-3):
    """the oracle's train_epoch / eval_epoch with the draw sequence of the product loop replayed from the current seeds:
    per minibatch G python draws of t (random.choice, :56), then one [N,3] and one [N,A] device normal_()"""
    p = {k: v.clone().requires_grad_(train) for k, v in sd.items()}
    opt = torch.optim.Adam(list(p.values()), lr=lr) if train else None
    fwd = lambda e, h, x: egnn_ref.egnn_forward(p, sizes["L"], e, h, x)
    epoch_loss, total_nodes = 0.0, 0
    for b in loader:
        Gn = b.num_graphs
        time_list = list(range(1, T + 1))
        t = torch.tensor([random.choice(time_list) for _ in range(Gn)])
        N = b.pos.shape[0]
        n_pos = torch.empty(N, 3, device=DEV).normal_(mean=0, std=1).cpu()
        n_h = torch.empty(N, A, device=DEV).normal_(mean=0, std=1).cpu()
        with torch.set_grad_enabled(train):
            loss, _ = LP.train_loss(fwd, sch, b.pos, b.x, b.edge_index, b.batch, t, n_pos, n_h, A=A)
        if train:
            opt.zero_grad()
            loss.backward()
            opt.step()
        epoch_loss += float(loss.detach()) * Gn   # (:178)
        total_nodes += N
    return epoch_loss / total_nodes, p             # (:181)


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_eval_epoch_and_train_epoch_return_values_match_oracle(lib, precision):
    from diffusion_model_b200.data import DataLoader
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.optim import FlatAdam
    from diffusion_model_b200.train_per_iretation import eval_epoch, train_epoch
    T = 1000
    net, sd, sizes = _tamed_net(3, 64, seed=2)
    net.precision = precision
    dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=T)
    sch = D.Schedule(1e-5, 2.0, T)
    loader = DataLoader(_dataset(7), batch_size=3)
    prm = dict(atom_type_size=A)
    opt = FlatAdam(net, lr=1e-3)
    # eval_epoch: no parameter update
    random.seed(9); torch.manual_seed(9)
    got = eval_epoch({"egnn": net}, loader, prm, dp, opt)
    random.seed(9); torch.manual_seed(9)
    want, _ = _replay_epoch(sd, sizes, loader, sch, T, train=False)
    assert abs(got - want) <= 2e-5 * abs(want), (got, want)
    assert all(torch.equal(p.detach().cpu(), sd[k]) for k, p in net.named_parameters())  # eval leaves the weights alone
    # train_epoch: three Adam steps inside the epoch; the returned per-node loss and the final weights
    random.seed(10); torch.manual_seed(10)
    got = train_epoch({"egnn": net}, loader, prm, dp, opt)
    random.seed(10); torch.manual_seed(10)
    want, p_ref = _replay_epoch(sd, sizes, loader, sch, T, train=True)
    assert abs(got - want) <= 1e-4 * abs(want), (got, want)
    # 3 Adam steps of lr 1e-3.  Adam's first steps are sign-like (update = lr g / (|g| + eps)): an entry whose gradient is
    # ~0 can land on the other side under a 1e-5 relative change of the gradients, i.e. 2 lr apart per step.  So: the bulk
    # within 2e-4, stragglers rare (< 0.5 %) and never further than the 3 x 2 lr a sign flip per step can produce
    diffs = torch.cat([(p.detach().cpu() - p_ref[k].detach()).abs().flatten() for k, p in net.named_parameters()])
    assert float((diffs > 2e-4).float().mean()) <= 5e-3, float((diffs > 2e-4).float().mean())
    assert float(diffs.max()) <= 6.1e-3, float(diffs.max())


def test_diffusion_process_loss_and_sample_aliases(lib):
    """dp.loss() == the minibatch loss of train_epoch, dp.sample() == generate()'s batched trajectory (BASELINE.json wording)"""
    from diffusion_model_b200.data import Batch
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    T = 30
    net, sd, sizes = _tamed_net(2, 64, seed=3)
    dp = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=T)
    sch = D.Schedule(0.05, 2.0, T)
    b = Batch.from_data_list(_dataset(3))
    gen = torch.Generator().manual_seed(4)
    t = torch.randint(1, T + 1, (3,), generator=gen)
    N = b.pos.shape[0]
    n_pos, n_h = torch.randn(N, 3, generator=gen), torch.randn(N, A, generator=gen)
    pr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    want, _ = LP.train_loss(lambda e, h, x: egnn_ref.egnn_forward(pr, sizes["L"], e, h, x), sch, b.pos, b.x, b.edge_index,
                            b.batch, t, n_pos, n_h, A=A)
    want.backward()
    bd = Batch.from_data_list(_dataset(3)).to(DEV)
    loss = dp.loss(net, bd, dict(atom_type_size=A), t=t.tolist(), noise_pos=n_pos.to(DEV), noise_h=n_h.to(DEV))
    assert loss.requires_grad and abs(loss.item() - want.item()) <= 1e-5 * abs(want.item())
    loss.backward()
    for k, p in net.named_parameters():
        assert relerr(p.grad, pr[k].grad) <= 1e-3, k
    # sample(): S structures in one call, injected noise, against the oracle's one-structure-at-a-time loop
    n, S = 12, 4
    per = [LP.make_noises(n, T, A, gen) for _ in range(S)]
    cat = {k: torch.cat([q[k] for q in per], dim=1 if k.startswith("step") else 0) for k in per[0]}
    out = dp.sample(net, n, S, noises=cat)
    ei = torch.as_tensor(G.fc_edges(n)).long()
    fwd = lambda e, h, x: egnn_ref.egnn_forward(sd, sizes["L"], e, h, x)
    for q in range(S):
        with torch.no_grad():
            pos, hf, atoms, finite = LP.sample_structure(fwd, sch, ei, n, per[q])
        assert finite and bool(out["finite"][q])
        assert relerr(out["pos"][q], pos) <= 2e-4 and relerr(out["h"][q], hf) <= 2e-4
        assert torch.equal(out["atoms"][q].cpu(), atoms)


@pytest.mark.parametrize("use_cuda_graph", [True, False])
def test_sampler_noise_ring_refills_match_oracle(lib, use_cuda_graph):
    """The sampler draws its own noise into a ring of `noise_chunk` steps that is refilled between CUDA-graph replays.
    With T = 40 and a 16-step ring the trajectory crosses three refills; replaying the SAME device generator in the same
    draw order and feeding those draws to the oracle's one-structure loop must reproduce every structure -- a deterministic
    check of the noise plumbing (ring indexing, refill timing, graph replay), where a distribution test can only hint."""
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.sampler import sample_structures
    T, n, S, chunk = 40, 10, 3, 16
    net, sd, sizes = _tamed_net(2, 64, seed=6)
    dp = E3DiffusionProcess(s=0.05, power=2.0, num_diffusion_timestep=T)
    sch = D.Schedule(0.05, 2.0, T)
    g = torch.Generator(device=DEV)
    g.manual_seed(77)
    with torch.no_grad():
        out = sample_structures(net, dp, n, S, generator=g, noise_chunk=chunk, use_cuda_graph=use_cuda_graph)
    g.manual_seed(77)
    N = S * n
    r = lambda *shape: torch.randn(*shape, device=DEV, generator=g)
    fill = lambda: (torch.empty(chunk, N, 3, device=DEV).normal_(generator=g), torch.empty(chunk, N, A, device=DEV).normal_(generator=g))
    init_x, init_h = r(N, 3), r(N, A)
    if use_cuda_graph:
        fill()  # the warm-up step before capture consumes one fill
    fills = [fill() for _ in range((T + chunk - 1) // chunk)]
    step_x = torch.cat([f[0] for f in fills])[:T]
    step_h = torch.cat([f[1] for f in fills])[:T]
    final_x, final_h = r(N, 3), r(N, A)
    ei = torch.as_tensor(G.fc_edges(n)).long()
    fwd = lambda e, h, x: egnn_ref.egnn_forward(sd, sizes["L"], e, h, x)
    for q in range(S):
        sl = slice(q * n, (q + 1) * n)
        noises = dict(init_x=init_x[sl].cpu(), init_h=init_h[sl].cpu(), step_x=step_x[:, sl].cpu(), step_h=step_h[:, sl].cpu(),
                      final_x=final_x[sl].cpu(), final_h=final_h[sl].cpu())
        with torch.no_grad():
            pos, hf, atoms, finite = LP.sample_structure(fwd, sch, ei, n, noises)
        assert finite and bool(out["finite"][q])
        assert relerr(out["pos"][q], pos) <= 2e-4 and relerr(out["h"][q], hf) <= 2e-4, (q, relerr(out["pos"][q], pos))
        assert torch.equal(out["atoms"][q].cpu(), atoms)


@pytest.mark.parametrize("precision", ["fp32", "bf16x3"])
def test_graphed_train_step_equals_eager_steps(lib, precision):
    """train_per_iretation.GraphedTrainStep (one CUDA-graph replay per minibatch: edge-list preparation, diffusion, forward,
    loss, backward, capturable fused Adam with its step counter on the device) against the same minibatches run eagerly,
    same t and noise: identical losses, weights equal to the rounding of the bias-correction powers; and constructing the
    graphed step (its warm-up and capture run real steps) must leave weights and optimizer state untouched."""
    from diffusion_model_b200.data import Batch
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.optim import FlatAdam
    from diffusion_model_b200.train_per_iretation import GraphedTrainStep, batch_loss
    T = 1000
    dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=T)
    prm = dict(atom_type_size=A)
    net_e, sd, sizes = _tamed_net(3, 64, seed=8)
    net_g, _, _ = _tamed_net(3, 64, seed=8)
    net_e.precision = net_g.precision = precision
    data = _dataset(3, seed=12)
    other = _dataset(3, seed=13)   # same shapes (sizes depend on the index only), different content
    b0 = Batch.from_data_list(data)
    opt_e, opt_g = FlatAdam(net_e, lr=1e-3), FlatAdam(net_g, lr=1e-3, capturable=True)
    w0 = net_g.flat_parameters().clone()
    gs = GraphedTrainStep(net_g, dp, prm, opt_g, b0, static_noise=True)
    assert torch.equal(net_g.flat_parameters(), w0) and float(opt_g.exp_avg.abs().max()) == 0.0
    gen = torch.Generator().manual_seed(3)
    N = b0.pos.shape[0]
    for it in range(4):
        b = Batch.from_data_list(data if it % 2 == 0 else other)
        t = torch.randint(1, T + 1, (3,), generator=gen).tolist()
        n_pos, n_h = torch.randn(N, 3, generator=gen).to(DEV), torch.randn(N, A, generator=gen).to(DEV)
        opt_e.zero_grad()
        loss_e, _, _ = batch_loss(net_e, Batch.from_data_list(data if it % 2 == 0 else other).to(DEV), dp, prm, t=t,
                                  noise_pos=n_pos, noise_h=n_h)
        loss_e.backward()
        opt_e.step()
        gs.noise_pos.copy_(n_pos)
        gs.noise_h.copy_(n_h)
        loss_g = gs.step(b, t=t)       # host minibatch -> staged copy -> static buffers -> replay
        assert float(loss_g) == float(loss_e) or abs(float(loss_g) - float(loss_e)) <= 1e-6 * abs(float(loss_e)), (it, float(loss_g), float(loss_e))
        gs.check()
        d = float((net_g.flat_parameters() - net_e.flat_parameters()).abs().max())
        assert d <= 2e-6, (it, d)      # lr 1e-3 x relative rounding of 1 - beta^t (device powf vs host powf)
    assert opt_g.state_dict()["step"] == 4
    with pytest.raises(Exception):     # a minibatch of another shape cannot go through the captured step
        gs.step(Batch.from_data_list(_dataset(2, seed=1)))
