"""Batched T-step ancestral sampler (the hot loop of generate(), parts/train_per_iretation.py:299-428).

S independent structures of n atoms are one batch of S graphs: every step is ONE denoiser call over all of them
(normaliser per structure, ``norm_scope='graph'`` == the reference's one-structure-per-call arithmetic) followed by ONE
fused kernel doing eps = P(x'-z), the posterior mean, the COM-free noise, the h update, the next time column and the
per-structure finite flag.  The per-step launch sequence is captured once in a CUDA graph and replayed for
t = T..1 with t held in a device counter; noise is drawn by torch in chunks into a ring the kernel indexes by t.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import call, ptr, stream_ptr
from .graph import GraphIndex, build_edges_fc

I32 = torch.int32


def sample_structures(egnn, dp, num_atoms, num_structures=1, atom_type_size=2, onehot_scale=1.0, cond=None,
                      edge_index=None, noises=None, generator=None, use_cuda_graph=True, noise_chunk=50,
                      record_steps=(), device="cuda", norm_scope="graph"):
    """Returns dict(pos [S,n,3], h [S,n,A], atoms [S,n,A] one-hot argmax, finite [S] bool, edge_index_single).

    edge_index: int64 [2,e] edges of ONE structure (default: fully connected, row-major, :308-313); it is replicated
    with node offsets for the S structures.  noises: optional injected draws, a dict of tensors shaped
    init_x [S*n,3], init_h [S*n,A], step_x [T,S*n,3], step_h [T,S*n,A] (row k is used at t = T-k), final_x, final_h --
    the layout of oracle/loops_ref.make_noises stacked over structures; otherwise torch.randn on the device.
    cond: optional [n,C] conditioning block inserted between the atom-type columns and the time column.
    norm_scope: 'graph' (default) = the normaliser of EGCL.message per structure, what generate() computes one structure
    at a time; 'call' = ONE scalar over the whole batch, what a batched reference call would compute (:64).
    """
    dev = torch.device(device)
    S, n, A = int(num_structures), int(num_atoms), int(atom_type_size)
    N = S * n
    T = dp.num_diffusion_timestep
    net = egnn._net
    Cw = 0 if cond is None else cond.shape[1]
    H = A + Cw + 1
    if net.H != H:
        raise _lib.EgnnError(f"network h width {net.H} != atom types {A} + conditioning {Cw} + 1")
    if edge_index is None:
        single = build_edges_fc([n], dev)
        ei = build_edges_fc([n] * S, dev)
    else:
        single = edge_index.to(dev)
        offs = (torch.arange(S, device=dev) * n).view(S, 1, 1)
        ei = (single.view(1, 2, -1) + offs).permute(1, 0, 2).reshape(2, -1).contiguous()
    node_ptr = (torch.arange(S + 1, device=dev) * n).to(I32)
    graph = GraphIndex(ei, N, node_ptr=node_ptr)
    _, _, coef = dp.device_tables(dev)

    def randn(*shape):
        return torch.randn(*shape, device=dev, generator=generator)

    inj = noises is not None
    z_x = (noises["init_x"].to(dev).float().clone() if inj else randn(N, 3)).contiguous()
    call("egnn_remove_mean", ptr(z_x), 3, ptr(node_ptr), S, stream_ptr(dev))  # (:303)
    z_h = (noises["init_h"].to(dev).float() if inj else randn(N, A))
    h_in = torch.empty(N, H, device=dev)
    h_in[:, :A] = onehot_scale * z_h
    if Cw:
        h_in[:, A:A + Cw] = cond.to(dev).float().repeat(S, 1)
    h_in[:, H - 1] = torch.tensor(T / T, dtype=torch.float32)
    h_out = torch.empty(N, H, device=dev)
    x_out = torch.empty(N, 3, device=dev)
    finite = torch.ones(S, dtype=I32, device=dev)
    t_dev = torch.full((1,), T, dtype=I32, device=dev)
    z_h_out = torch.empty(N, A, device=dev)

    if inj:
        ring_x = noises["step_x"].to(dev).float().contiguous()
        ring_h = noises["step_h"].to(dev).float().contiguous()
        period, chunk = 0, T
    else:
        chunk = max(1, min(int(noise_chunk), T))
        ring_x = torch.empty(chunk, N, 3, device=dev)
        ring_h = torch.empty(chunk, N, A, device=dev)
        period = chunk

    def one_step():
        egnn.forward_into(graph, h_in, z_x, h_out, x_out, norm_scope=norm_scope)
        call("egnn_sampler_step", ptr(z_x), ptr(h_in), ptr(x_out), ptr(h_out), ptr(ring_x), ptr(ring_h), ptr(coef),
             ptr(t_dev), T, N * 3, N * A, period, float(onehot_scale), ptr(node_ptr), S, H, A, 0, None, ptr(finite),
             stream_ptr(dev))
        call("egnn_counter_add", ptr(t_dev), -1, stream_ptr(dev))

    cuda_graph = None
    if use_cuda_graph and not record_steps:
        # warm-up on a side stream (allocates the workspace, packs weights), then capture; the warm-up steps are real
        # work on scratch copies so that the captured step starts from the true state
        keep = (z_x.clone(), h_in.clone(), finite.clone())
        if not inj:
            ring_x.normal_(generator=generator)
            ring_h.normal_(generator=generator)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            one_step()
        torch.cuda.current_stream(dev).wait_stream(side)
        z_x.copy_(keep[0]); h_in.copy_(keep[1]); finite.copy_(keep[2]); t_dev.fill_(T)
        cuda_graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(cuda_graph):
            one_step()
        z_x.copy_(keep[0]); h_in.copy_(keep[1]); finite.copy_(keep[2]); t_dev.fill_(T)

    record = {}
    for t in range(T, 0, -1):
        if not inj and (T - t) % chunk == 0:
            ring_x.normal_(generator=generator)
            ring_h.normal_(generator=generator)
        if cuda_graph is not None:
            cuda_graph.replay()
        else:
            one_step()
        if t in record_steps:
            record[t] = (z_x.clone(), h_in[:, :A].clone() / onehot_scale)
    # final t = 0 step (:391-428)
    h_in[:, H - 1] = 0.0
    egnn.forward_into(graph, h_in, z_x, h_out, x_out, norm_scope=norm_scope)
    fx = (noises["final_x"].to(dev).float() if inj else randn(N, 3)).contiguous()
    fh = (noises["final_h"].to(dev).float() if inj else randn(N, A)).contiguous()
    call("egnn_sampler_step", ptr(z_x), ptr(h_in), ptr(x_out), ptr(h_out), ptr(fx), ptr(fh), ptr(coef), None, 0, 0, 0, 0,
         float(onehot_scale), ptr(node_ptr), S, H, A, 1, ptr(z_h_out), ptr(finite), stream_ptr(dev))
    atoms = torch.nn.functional.one_hot(torch.argmax(z_h_out, dim=1), num_classes=A)
    return dict(pos=z_x.view(S, n, 3), h=z_h_out.view(S, n, A), atoms=atoms.view(S, n, A), finite=finite.bool(),
                edge_index_single=single, record=record, graph=graph)
