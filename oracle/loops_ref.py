"""TEST INFRASTRUCTURE ONLY (CPU oracle) -- never imported by the product package.

Restated training-loss and sampler loops with INJECTABLE diffusion times and noise.  Follows:
  * parts/train_per_iretation.py:36-92   diffuse_as_batch: per graph t (python random.choice there; an argument
                                         here), pos noise COM-free per graph, atom-type noise plain
  * parts/train_per_iretation.py:122-181 train_epoch body: time/T, h=[h_t | t/T], egnn, eps_x=P(x'-pos_t),
                                         eps_h=h'[:, :A], loss = SSE / num_graph
  * parts/train_per_iretation.py:299-442 generate body: T ancestral steps + the t=0 step (:391-428), argmax
  * train.py:238-288 / test.py:226-279   x-only loop (target = position noise only)
The EGNN itself is a callable ``egnn(edge_index, h, x) -> (h, x)`` so the same loops can drive the restatement
(oracle/egnn_ref.py), the unmodified reference module (build container only) or anything else under test.
Unconditional model only (h = [atom types | t/T], H = A + 1); the conditional concatenations
(:136-145, :340-352) are represented by the optional ``cond`` [N,C] block inserted before the time column.
"""
from __future__ import annotations

import torch

from . import diffusion_ref as D


def time_feature(t, T, n):
    """[n,1] fp32 column t/T -- fp32(t)/fp32(T) as train_epoch:133 computes it (long tensor / int)."""
    return (torch.full((n, 1), int(t), dtype=torch.long) / T).to(torch.float32)


def diffuse_batch(sch, pos, x_onehot, batch, t_per_graph, noise_pos, noise_h):
    """diffuse_as_batch with injected draws.  Returns pos_t, h_t, eps_pos (COM-free per graph), eps_h, t_node."""
    G = int(batch.max()) + 1
    pos_t, h_t, e_pos, e_h, t_node = [], [], [], [], []
    for g in range(G):
        sel = batch == g
        t = int(t_per_graph[g])
        p, n_p = D.diffuse_zero_to_t(sch, pos[sel], t, noise_pos[sel].clone(), "pos")
        hh, n_h = D.diffuse_zero_to_t(sch, x_onehot[sel], t, noise_h[sel].clone(), "h")
        pos_t.append(p); h_t.append(hh); e_pos.append(n_p); e_h.append(n_h)
        t_node.append(torch.full((int(sel.sum()), 1), t, dtype=torch.long))
    return torch.cat(pos_t), torch.cat(h_t), torch.cat(e_pos), torch.cat(e_h), torch.cat(t_node)


def train_loss(egnn, sch, pos, x_onehot, edge_index, batch, t_per_graph, noise_pos, noise_h, A=2, cond=None):
    """One minibatch of train_epoch up to (and including) ``loss = loss / num_graph``.  Returns (loss, aux)."""
    G = int(batch.max()) + 1
    pos_t, h_t, e_pos, e_h, t_node = diffuse_batch(sch, pos, x_onehot, batch, t_per_graph, noise_pos, noise_h)
    time_data = t_node / sch.T
    parts = [h_t] + ([cond] if cond is not None else []) + [time_data]
    h_in = torch.cat(parts, dim=1)
    h, x = egnn(edge_index, h_in, pos_t)
    eps_x = D.remove_mean((x - pos_t), batch)
    eps_h = h[:, :A]
    pred = torch.cat((eps_x, eps_h), dim=1)
    target = torch.cat((e_pos, e_h), dim=1)
    loss = ((pred - target) ** 2).sum() / G
    return loss, dict(pos_t=pos_t, h_t=h_t, eps_pos=e_pos, eps_h=e_h, h_in=h_in, h_out=h, x_out=x)


def sample_structure(egnn, sch, edge_index, n, noises, A=2, s=1.0, cond=None, record=None):
    """generate()'s inner loop for ONE structure with injected noise.

    noises: dict(init_x [n,3], init_h [n,A], step_x [T,n,3], step_h [T,n,A] (index T-t), final_x, final_h).
    Returns (pos [n,3], h [n,A], atoms one-hot [n,2], finite flag).
    """
    T = sch.T
    z_x = D.remove_mean(noises["init_x"].clone())
    z_h = noises["init_h"].clone()
    finite = True
    for t in range(T, 0, -1):
        k = T - t
        tcol = torch.full((n, 1), t / T, dtype=torch.float32)  # generate():340 -- python float then fp32
        parts = [s * z_h] + ([cond] if cond is not None else []) + [tcol]
        h = torch.cat(parts, dim=1)
        new_h, new_x = egnn(edge_index, h, z_x)
        eps_x = D.remove_mean(new_x - z_x)
        eps_h = new_h[:, :A]
        z_x = D.reverse_diffuse_one_step(sch, z_x, eps_x, t, noises["step_x"][k].clone(), "pos")
        z_h = D.reverse_diffuse_one_step(sch, h[:, :A], eps_h, t, noises["step_h"][k].clone(), "h")
        if record is not None:
            record(t, z_x, z_h)
        if not (torch.isfinite(z_h).all() and torch.isfinite(z_x).all()):
            finite = False
            break
    tcol = torch.zeros((n, 1), dtype=torch.float32)
    parts = [s * z_h] + ([cond] if cond is not None else []) + [tcol]
    h = torch.cat(parts, dim=1)
    new_h, new_x = egnn(edge_index, h, z_x)
    eps_x = D.remove_mean(new_x - z_x)
    hh = h[:, :A]
    eps_h = new_h[:, :A]
    a0, s0 = sch.alpha[0], sch.sigma[0]
    pos = z_x / a0 - s0 * eps_x / a0 + s0 * D.remove_mean(noises["final_x"].clone()) / a0
    hf = hh / a0 - s0 * eps_h / a0 + s0 * noises["final_h"] / a0
    atoms = torch.nn.functional.one_hot(torch.argmax(hf, dim=1), num_classes=2)
    finite = finite and bool(torch.isfinite(pos).all()) and bool(torch.isfinite(hf).all())
    return pos, hf, atoms, finite


def make_noises(n, T, A, generator):
    r = lambda *s: torch.randn(*s, generator=generator)
    return dict(init_x=r(n, 3), init_h=r(n, A), step_x=r(T, n, 3), step_h=r(T, n, A), final_x=r(n, 3), final_h=r(n, A))


# ---- x-only loop (train.py:238-288, test.py:253-279) ----------------------------------------------------
def train_loss_xonly(egnn, sch, pos, x_onehot, edge_index, t, noise_pos, s=1.0):
    """train.py's single-graph step: h=[s*onehot | t/T], target = COM-free position noise, loss = SSE."""
    n = pos.shape[0]
    pos_t, eps = D.diffuse_zero_to_t(sch, pos, t, noise_pos.clone(), "pos")
    h = torch.cat((s * x_onehot, time_feature(t, sch.T, n)), dim=1)
    _, x = egnn(edge_index, h, pos_t)
    eps_x = D.remove_mean(x - pos_t)
    return ((eps_x - eps) ** 2).sum()


def sample_structure_xonly(egnn, sch, edge_index, x_onehot, noises, s=1.0):
    T, n = sch.T, x_onehot.shape[0]
    pos = D.remove_mean(noises["init_x"].clone())
    for t in range(T, 0, -1):
        h = torch.cat((s * x_onehot, torch.full((n, 1), t / T, dtype=torch.float32)), dim=1)
        _, new_x = egnn(edge_index, h, pos)
        eps = D.remove_mean(new_x - pos)
        mu = D.calculate_mu_xonly(sch, pos, eps, t)
        pos = D.reverse_diffuse_one_step_xonly(sch, mu, t, noises["step_x"][T - t].clone())
    return pos


def sample_structures_batched(sd, L, sch, n, S, noises, A=2, s=1.0, norm_scope="graph"):
    """S independent runs of sample_structure() evaluated as ONE batch of S graphs (faster on the CPU, same arithmetic
    per structure): per-graph COM removal (diffusion_x_h.py:10-13 with a batch vector) and the per-graph normaliser
    (norm_scope='graph'), which is what calling generate() once per structure (parts/train_per_iretation.py:299-316)
    computes.  noises: dict as make_noises() with a leading structure axis flattened into the node axis
    (init_x [S*n,3], step_x [T,S*n,3], ...).  Returns (pos [S,n,3], h [S,n,A], finite [S])."""
    from . import egnn_ref, graphs_ref
    T = sch.T
    batch = torch.arange(S).repeat_interleave(n)
    ei = torch.as_tensor(graphs_ref.fc_edges_batch([n] * S))
    # norm_scope='call': ONE normaliser over the whole batch -- what EGCL.message (:64) computes for a batched call
    fwd = lambda h, x: egnn_ref.egnn_forward(sd, L, ei, h, x, norm_scope, batch=batch)

    def rm(x, _b=None):  # per-graph COM removal; equal-size graphs -> one strided mean instead of the mask loop
        v = x.view(S, n, 3)
        return (v - v.mean(dim=1, keepdim=True)).reshape(S * n, 3)
    z_x = rm(noises["init_x"].clone())
    z_h = noises["init_h"].clone()
    for t in range(T, 0, -1):
        k = T - t
        h = torch.cat([s * z_h, torch.full((S * n, 1), t / T, dtype=torch.float32)], dim=1)
        new_h, new_x = fwd(h, z_x)
        eps_x = rm(new_x - z_x)
        z_x = D.reverse_diffuse_one_step(sch, z_x, eps_x, t, rm(noises["step_x"][k].clone()), "h")
        z_h = D.reverse_diffuse_one_step(sch, h[:, :A], new_h[:, :A], t, noises["step_h"][k].clone(), "h")
    h = torch.cat([s * z_h, torch.zeros((S * n, 1), dtype=torch.float32)], dim=1)
    new_h, new_x = fwd(h, z_x)
    eps_x = rm(new_x - z_x)
    a0, s0 = sch.alpha[0], sch.sigma[0]
    pos = z_x / a0 - s0 * eps_x / a0 + s0 * rm(noises["final_x"].clone()) / a0
    hf = h[:, :A] / a0 - s0 * new_h[:, :A] / a0 + s0 * noises["final_h"] / a0
    pos, hf = pos.view(S, n, 3), hf.view(S, n, A)
    finite = torch.isfinite(pos).all(dim=(1, 2)) & torch.isfinite(hf).all(dim=(1, 2))
    return pos, hf, finite
