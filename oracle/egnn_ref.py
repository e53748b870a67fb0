"""TEST INFRASTRUCTURE ONLY (CPU oracle) -- never imported by the product package.

Independent pure-torch CPU restatement of the reference EGNN denoiser.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import it.

Follows (file:line under the reference checkout):
  * EquivariantGraphNeuralNetwork.py:13-34  -- the four MLPs of one EGCL (Linear/SiLU stacks, sigmoid gate)
  * EquivariantGraphNeuralNetwork.py:55-65  -- message(): u=[h_i,h_j,||x_i-x_j||^2]; gated m; coordinate message
                                               (x_i-x_j)*phi/(||X_i-X_j||_F+1) with a SINGLE global norm
  * EquivariantGraphNeuralNetwork.py:67-71  -- forward(): sum aggregation on edge_index[0]
                                               (flow='target_to_source', :10-11), h'=mlp_h([h,M]) (no residual),
                                               x'=x+sum
  * EquivariantGraphNeuralNetwork.py:85-88  -- L layers applied sequentially
torch_geometric's contribution (gather by edge_index rows, scatter-add on edge_index[0]) is restated
with index_select / index_add_ (a second, independent formulation from oracle/pyg_stub.py).

Parity pin: the reference ships no tests, fixtures or golden vectors for this path, so the
pin is "outputs of the reference itself run in the build container": ``oracle/make_golden.py`` imports the
unmodified reference file through the stub and asserts this restatement reproduces it to fp32 round-off,
then freezes vectors under tests/golden/.

Parameter naming follows the reference ``state_dict``:
  egcl_list.{l}.mlp_m.{0,2}.{weight,bias}, .mlp_x.{0,2,4}., .mlp_h.{0,2}., .attention.0.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def silu(z):
    return z * torch.sigmoid(z)


def layer_params(state_dict, l):
    """Extract one EGCL's tensors from a reference-style state_dict."""
    p = f"egcl_list.{l}."
    g = lambda k: state_dict[p + k]
    return dict(
        wm1=g("mlp_m.0.weight"), bm1=g("mlp_m.0.bias"), wm2=g("mlp_m.2.weight"), bm2=g("mlp_m.2.bias"),
        wx1=g("mlp_x.0.weight"), bx1=g("mlp_x.0.bias"), wx2=g("mlp_x.2.weight"), bx2=g("mlp_x.2.bias"),
        wx3=g("mlp_x.4.weight"), bx3=g("mlp_x.4.bias"),
        wh1=g("mlp_h.0.weight"), bh1=g("mlp_h.0.bias"), wh2=g("mlp_h.2.weight"), bh2=g("mlp_h.2.bias"),
        wa=g("attention.0.weight"), ba=g("attention.0.bias"),
    )


def edge_frobenius(x, edge_index, norm_scope="call", edge_graph=None, num_graphs=None):
    """S of EquivariantGraphNeuralNetwork.py:64 -- torch.norm(coords_i-coords_j) with no dim.

    norm_scope='call'  : one scalar over every edge of the call (what a batched reference call computes).
    norm_scope='graph' : one scalar per graph (what the reference computes when it samples structures one
                         at a time, parts/train_per_iretation.py:299-316); returns a per-edge tensor [E,1].
    """
    r = x.index_select(0, edge_index[0]) - x.index_select(0, edge_index[1])
    if norm_scope == "call":
        return torch.sqrt((r * r).sum())
    sq = (r * r).sum(dim=1)
    per_graph = torch.zeros(num_graphs, dtype=x.dtype).index_add_(0, edge_graph, sq)
    return torch.sqrt(per_graph).index_select(0, edge_graph).unsqueeze(1)


def egcl_forward(p, edge_index, h, x, norm_scope="call", edge_graph=None, num_graphs=None):
    """One EGCL.  p: dict from layer_params().  Returns (h', x')."""
    i, j = edge_index[0], edge_index[1]
    h_i, h_j = h.index_select(0, i), h.index_select(0, j)
    r = x.index_select(0, i) - x.index_select(0, j)
    # :56 -- sqrt then square, kept on purpose
    d = torch.sqrt((r * r).sum(dim=1, keepdim=True)) ** 2
    u = torch.cat((h_i, h_j, d), dim=1)
    # mode 'h' (:57-61)
    m = silu(F.linear(silu(F.linear(u, p["wm1"], p["bm1"])), p["wm2"], p["bm2"]))
    m = m * torch.sigmoid(F.linear(m, p["wa"], p["ba"]))
    M = torch.zeros(h.size(0), m.size(1), dtype=h.dtype).index_add_(0, i, m)
    h_new = F.linear(silu(F.linear(torch.cat((h, M), dim=1), p["wh1"], p["bh1"])), p["wh2"], p["bh2"])
    # mode 'x' (:62-65)
    phi = F.linear(silu(F.linear(silu(F.linear(u, p["wx1"], p["bx1"])), p["wx2"], p["bx2"])), p["wx3"], p["bx3"])
    S = edge_frobenius(x, edge_index, norm_scope, edge_graph, num_graphs)
    out = r * phi / (S + 1)
    x_new = x + torch.zeros_like(x).index_add_(0, i, out)
    return h_new, x_new


def egnn_forward(state_dict, L, edge_index, h, x, norm_scope="call", batch=None, return_layers=False, checkpoint=False):
    """EquivariantGNN.forward(edge_index, h, x) -> (h, x)  (EquivariantGraphNeuralNetwork.py:85-88).

    checkpoint=True recomputes each layer in backward (same arithmetic, ~L times less autograd memory: the full-size
    parity tests run E = 4e5 edges through 6-9 layers on the host)."""
    edge_graph = num_graphs = None
    if norm_scope == "graph":
        if batch is None:
            batch = torch.zeros(h.size(0), dtype=torch.long)
        edge_graph = batch.index_select(0, edge_index[0])
        num_graphs = int(batch.max()) + 1
    layers = []
    for l in range(L):
        if checkpoint:
            from torch.utils.checkpoint import checkpoint as _ckpt
            p = layer_params(state_dict, l)
            keys = list(p)
            h, x = _ckpt(lambda hh, xx, *w: egcl_forward(dict(zip(keys, w)), edge_index, hh, xx, norm_scope, edge_graph,
                                                         num_graphs), h, x, *[p[k] for k in keys], use_reentrant=False)
        else:
            h, x = egcl_forward(layer_params(state_dict, l), edge_index, h, x, norm_scope, edge_graph, num_graphs)
        if return_layers:
            layers.append((h, x))
    return (h, x, layers) if return_layers else (h, x)


def init_state_dict(L, m_input, m_hidden, m_output, x_input, x_hidden, x_output, h_input, h_hidden, h_output,
                    dtype=torch.float32):
    """Random weights with the reference's construction order and default nn.Linear init.

    The reference builds, per layer and in this order, mlp_m (2 Linear), mlp_x (3 Linear), mlp_h (2 Linear),
    attention (1 Linear) (EquivariantGraphNeuralNetwork.py:13-34), so drawing nn.Linear modules in the same
    order under the same torch seed reproduces ``EquivariantGNN(...)``'s initial weights exactly.
    """
    sd = {}
    for l in range(L):
        spec = [
            ("mlp_m.0", m_input, m_hidden), ("mlp_m.2", m_hidden, m_output),
            ("mlp_x.0", x_input, x_hidden), ("mlp_x.2", x_hidden, x_hidden), ("mlp_x.4", x_hidden, x_output),
            ("mlp_h.0", h_input, h_hidden), ("mlp_h.2", h_hidden, h_output),
            ("attention.0", m_output, 1),
        ]
        for name, fin, fout in spec:
            lin = torch.nn.Linear(fin, fout)
            sd[f"egcl_list.{l}.{name}.weight"] = lin.weight.detach().to(dtype).clone()
            sd[f"egcl_list.{l}.{name}.bias"] = lin.bias.detach().to(dtype).clone()
    return sd


def net_sizes(L, d, H=3, A=2):
    """Constructor arguments for a 'hidden d' network of SURVEY.md §8 (Mh=Ms=Xh=Hh=d, K0=2H+1)."""
    K0 = 2 * H + 1
    return dict(L=L, m_input=K0, m_hidden=d, m_output=d, x_input=K0, x_hidden=d, x_output=1,
                h_input=H + d, h_hidden=d, h_output=H)


def tame_state_dict(sd, L, fx=0.05, fd=0.02):
    """Deterministic down-scaling of random-init weights so that an UNTRAINED net gives a bounded T-step trajectory.

    No trained weights ship with the reference.  With default init the ancestral sampler diverges
    within 3 steps (z grows like alpha_t/alpha_T and ||x_i-x_j||^2 feeds the MLPs), so trajectory-level tests and the
    sampling bench scale the coordinate head (mlp_x.4) by fx and the distance column of both first layers by fd.
    The code path and shapes are unchanged.
    """
    sd = {k: v.clone() for k, v in sd.items()}
    for l in range(L):
        p = f"egcl_list.{l}."
        sd[p + "mlp_x.4.weight"] *= fx
        sd[p + "mlp_x.4.bias"] *= fx
        sd[p + "mlp_m.0.weight"][:, -1] *= fd
        sd[p + "mlp_x.0.weight"][:, -1] *= fd
    return sd
