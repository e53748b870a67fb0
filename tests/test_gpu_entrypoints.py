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
    dp = E3DiffusionProcess(s=1e-3, power=2.0, num_diffusion_timestep=T)
    sch = D.Schedule(1e-3, 2.0, T)
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
    for k, p in net.named_parameters():
        assert float((p.detach().cpu() - p_ref[k].detach()).abs().max()) <= 2e-4, k  # 3 steps of lr 1e-3 (sign-like Adam steps)


def test_diffusion_process_loss_and_sample_aliases(lib):
    """dp.loss() == the minibatch loss of train_epoch, dp.sample() == generate()'s batched trajectory (BASELINE.json wording)"""
    from diffusion_model_b200.data import Batch
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    T = 30
    net, sd, sizes = _tamed_net(2, 64, seed=3)
    dp = E3DiffusionProcess(s=1e-3, power=2.0, num_diffusion_timestep=T)
    sch = D.Schedule(1e-3, 2.0, T)
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
