"""GPU parity tests of the backward kernels and the training loss: gradients from the unmodified reference module's
autograd (golden 'g:' arrays, oracle/make_golden.py) and from the CPU oracle's autograd on the same seeded inputs.

Bars: loss 1e-5 relative; gradients 1e-3 relative per tensor (fixed-order vs edge-order summation), SURVEY.md 8(c).
"""
import os

import numpy as np
import pytest
import torch

from oracle import diffusion_ref as D
from oracle import egnn_ref, graphs_ref as G, loops_ref as LP

pytestmark = pytest.mark.gpu
DEV = "cuda"
GTOL = 1e-3


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def state_dict_of(z):
    return {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w:")}


def sizes_of(z):
    return {str(k): int(v) for k, v in zip(z["size_keys"], z["sizes"])}


def relerr(a, b):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def make_net(sizes, sd):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net.precision = "fp32"  # these tests pin the CUDA-core fp32 mode; the tcgen05 modes are pinned in test_gpu_tc.py
    return net.to(DEV)


def check_param_grads(net, ref_grads, tol=GTOL):
    worst = 0.0
    for k, p in net.named_parameters():
        assert p.grad is not None, k
        e = relerr(p.grad, ref_grads[k])
        worst = max(worst, e)
        assert e <= tol, (k, e)
    return worst


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg4", "ragged"])
def test_param_and_input_grads_match_reference(lib, golden_dir, name):
    z = load(golden_dir, "egnn_" + name)
    sizes, sd = sizes_of(z), state_dict_of(z)
    net = make_net(sizes, sd)
    h0, x0 = torch.from_numpy(z["h"]), torch.from_numpy(z["x"])
    ei = torch.from_numpy(z["edge_index"])
    th, tx = torch.from_numpy(z["target_h"]), torch.from_numpy(z["target_x"])
    h = h0.to(DEV).requires_grad_(True)
    x = x0.to(DEV).requires_grad_(True)
    ho, xo = net(ei.to(DEV), h, x)
    loss = ((ho - th.to(DEV)) ** 2).sum() + ((xo - x - tx.to(DEV)) ** 2).sum()
    loss.backward()
    assert abs(loss.item() - float(z["loss"])) <= 2e-5 * abs(float(z["loss"]))
    check_param_grads(net, {k[2:]: z[k] for k in z.files if k.startswith("g:")})
    # input gradients: oracle autograd on the CPU
    hc, xc = h0.clone().requires_grad_(True), x0.clone().requires_grad_(True)
    hr, xr = egnn_ref.egnn_forward(sd, sizes["L"], ei, hc, xc)
    (((hr - th) ** 2).sum() + ((xr - xc - tx) ** 2).sum()).backward()
    assert relerr(h.grad, hc.grad) <= GTOL and relerr(x.grad, xc.grad) <= GTOL
    # bit-identical reruns of the backward (no atomics)
    g1 = [p.grad.clone() for p in net.parameters()]
    net.zero_grad()
    h2, x2 = h0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
    ho, xo = net(ei.to(DEV), h2, x2)
    (((ho - th.to(DEV)) ** 2).sum() + ((xo - x2 - tx.to(DEV)) ** 2).sum()).backward()
    assert all(torch.equal(a, p.grad) for a, p in zip(g1, net.parameters()))
    assert torch.equal(h2.grad, h.grad) and torch.equal(x2.grad, x.grad)


def test_graph_scope_norm_backward(lib, golden_dir):
    z = load(golden_dir, "egnn_cfg1")
    sizes, sd = sizes_of(z), state_dict_of(z)
    net = make_net(sizes, sd)
    h0, x0, ei = torch.from_numpy(z["h"]), torch.from_numpy(z["x"]), torch.from_numpy(z["edge_index"])
    batch = torch.from_numpy(z["batch"])
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    hc, xc = h0.clone().requires_grad_(True), x0.clone().requires_grad_(True)
    hr, xr = egnn_ref.egnn_forward(sdg, sizes["L"], ei, hc, xc, "graph", batch=batch)
    (hr.square().sum() + (xr * xr.roll(1, 0)).sum()).backward()
    h, x = h0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
    ho, xo = net(ei.to(DEV), h, x, batch=batch.to(DEV), norm_scope="graph")
    (ho.square().sum() + (xo * xo.roll(1, 0)).sum()).backward()
    check_param_grads(net, {k: v.grad for k, v in sdg.items()})
    assert relerr(h.grad, hc.grad) <= GTOL and relerr(x.grad, xc.grad) <= GTOL


def test_backward_edge_cases(lib):
    """isolated nodes, a receiver cut by several tiles, an asymmetric edge list, and no edges at all"""
    torch.manual_seed(0)
    sizes = egnn_ref.net_sizes(2, 32)
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    net = EquivariantGNN(**sizes).to(DEV)
    net.precision = "fp32"
    sd = {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    gen = torch.Generator().manual_seed(1)
    n = 320
    recv = np.concatenate([np.full(300, 2), np.full(5, 7), np.full(200, 9)]).astype(np.int64)
    send = np.concatenate([np.arange(10, 310), np.arange(0, 5), np.arange(100, 300)]).astype(np.int64)
    for ei in (torch.from_numpy(np.stack([recv, send])), torch.zeros(2, 0, dtype=torch.int64)):
        h0, x0 = torch.randn(n, 3, generator=gen), torch.randn(n, 3, generator=gen)
        sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        hc, xc = h0.clone().requires_grad_(True), x0.clone().requires_grad_(True)
        hr, xr = egnn_ref.egnn_forward(sdg, 2, ei, hc, xc)
        (hr.square().sum() + xr.square().sum()).backward()
        net.zero_grad()
        h, x = h0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
        ho, xo = net(ei.to(DEV), h, x)
        (ho.square().sum() + xo.square().sum()).backward()
        for k, p in net.named_parameters():
            ref = sdg[k].grad if sdg[k].grad is not None else torch.zeros_like(sdg[k])
            scale = max(float(ref.abs().max()), 1e-3)
            assert float((p.grad.cpu() - ref).abs().max()) <= GTOL * scale, k
        assert relerr(h.grad, hc.grad) <= GTOL and relerr(x.grad, xc.grad) <= GTOL


def test_training_loss_and_grads_match_reference(lib, golden_dir):
    """the minibatch of parts/train_per_iretation.py:130-172 with injected t and noise (golden from the reference)"""
    from diffusion_model_b200.data import Data
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.train_per_iretation import batch_loss
    z = load(golden_dir, "loss_cfg1")
    sd = state_dict_of(z)
    net = make_net(egnn_ref.net_sizes(4, 64), sd)
    dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=1000)
    bd = Data(x=torch.from_numpy(z["x"]).to(DEV), pos=torch.from_numpy(z["pos"]).to(DEV),
              edge_index=torch.from_numpy(z["edge_index"]).to(DEV))
    bd.batch = torch.from_numpy(z["batch"]).to(DEV)
    loss, G_, aux = batch_loss(net, bd, dp, dict(atom_type_size=2), t=z["t"].tolist(),
                               noise_pos=torch.from_numpy(z["noise_pos"]).to(DEV),
                               noise_h=torch.from_numpy(z["noise_h"]).to(DEV))
    assert G_ == 4
    assert abs(loss.item() - float(z["loss"])) <= 1e-5 * abs(float(z["loss"]))
    assert relerr(aux["h_out"], z["h_out"]) <= 1e-4 and relerr(aux["x_out"], z["x_out"]) <= 1e-4
    loss.backward()
    check_param_grads(net, {k[2:]: z[k] for k in z.files if k.startswith("g:")})


def test_adam_kernel_and_training_steps(lib, golden_dir):
    """three optimiser steps: FlatAdam (fused kernel on the flat buffers) against torch.optim.Adam on the CPU oracle"""
    from diffusion_model_b200.data import Data
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.optim import FlatAdam
    from diffusion_model_b200.train_per_iretation import batch_loss
    z = load(golden_dir, "loss_cfg1")
    sd = state_dict_of(z)
    sizes = egnn_ref.net_sizes(4, 64)
    net = make_net(sizes, sd)
    opt = FlatAdam(net, lr=1e-3)
    dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=1000)
    sch = D.Schedule(1e-5, 2.0, 1000)
    ref_p = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    ref_opt = torch.optim.Adam(list(ref_p.values()), lr=1e-3)
    pos, x1h = torch.from_numpy(z["pos"]), torch.from_numpy(z["x"])
    ei, batch = torch.from_numpy(z["edge_index"]), torch.from_numpy(z["batch"])
    bd = Data(x=x1h.to(DEV), pos=pos.to(DEV), edge_index=ei.to(DEV))
    bd.batch = batch.to(DEV)
    gen = torch.Generator().manual_seed(77)
    for step in range(3):
        t = torch.randint(1, 1001, (4,), generator=gen)
        n_pos, n_h = torch.randn(96, 3, generator=gen), torch.randn(96, 2, generator=gen)
        ref_opt.zero_grad()
        lr_, _ = LP.train_loss(lambda e, h, x: egnn_ref.egnn_forward(ref_p, 4, e, h, x), sch, pos, x1h, ei, batch, t,
                               n_pos, n_h)
        lr_.backward()
        ref_opt.step()
        opt.zero_grad()
        loss, _, _ = batch_loss(net, bd, dp, dict(atom_type_size=2), t=t.tolist(), noise_pos=n_pos.to(DEV),
                                noise_h=n_h.to(DEV))
        loss.backward()
        opt.step()
        assert abs(loss.item() - lr_.item()) <= 1e-4 * abs(lr_.item()), (step, loss.item(), lr_.item())
    for k, p in net.named_parameters():
        assert float((p.detach().cpu() - ref_p[k].detach()).abs().max()) <= 2e-4, k  # 3 steps of lr 1e-3


def test_remove_mean_autograd(lib):
    from diffusion_model_b200.diffusion_x_h import remove_mean
    gen = torch.Generator().manual_seed(2)
    v = torch.randn(50, 3, generator=gen)
    w = torch.randn(50, 3, generator=gen)
    b = torch.repeat_interleave(torch.arange(5), 10)
    vc = v.clone().requires_grad_(True)
    (D.remove_mean(vc * 1.0, b) * w).sum().backward()
    vd = v.to(DEV).requires_grad_(True)
    (remove_mean(vd * 1.0, b.to(DEV)) * w.to(DEV)).sum().backward()
    assert relerr(vd.grad, vc.grad) < 1e-5


def test_device_prefetcher_and_train_epoch(lib, golden_dir):
    """DevicePrefetcher hands out exactly the loader's batches (copied on its side stream, one ahead), and train_epoch
    over host-resident data through it is reproducible: same seed -> bit-identical epoch loss and parameters."""
    import random
    from diffusion_model_b200.data import Data, DataLoader, DevicePrefetcher
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.optim import FlatAdam
    from diffusion_model_b200.train_per_iretation import train_epoch
    z = load(golden_dir, "loss_cfg1")
    sd = state_dict_of(z)
    gen = torch.Generator().manual_seed(5)
    data = []
    for g in range(10):
        n = 20 + 2 * g
        pos = 4.0 * torch.rand(n, 3, generator=gen)
        x = torch.nn.functional.one_hot(torch.randint(0, 2, (n,), generator=gen), 2).float()
        ei = torch.as_tensor(G.fc_edges(n)).long()
        data.append(Data(x=x, pos=pos, edge_index=ei))
    loader = DataLoader(data, batch_size=4)
    got = list(DevicePrefetcher(loader, DEV))
    want = list(loader)
    assert len(got) == len(want) == 3
    for a, b in zip(got, want):
        assert a.pos.is_cuda and a.edge_index.is_cuda and a.num_graphs == b.num_graphs
        for k in ("x", "pos", "edge_index", "batch", "node_ptr"):
            assert torch.equal(getattr(a, k).cpu(), getattr(b, k)), k
    assert list(DevicePrefetcher([], DEV)) == []

    def run():
        net = make_net(egnn_ref.net_sizes(4, 64), sd)
        opt = FlatAdam(net, lr=1e-3)
        dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=1000)
        random.seed(3)
        torch.manual_seed(3)
        losses = [train_epoch({"egnn": net}, loader, dict(atom_type_size=2), dp, opt) for _ in range(2)]
        return losses, torch.cat([p.detach().flatten() for p in net.parameters()]).cpu()

    l1, p1 = run()
    l2, p2 = run()
    assert l1 == l2 and all(np.isfinite(l1)) and torch.equal(p1, p2)
