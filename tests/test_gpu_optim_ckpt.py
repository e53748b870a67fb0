"""the schedule-free optimizer the reference's parameters.yaml selects (parts/def_for_main.py:137-138) as a
fused kernel on the flat buffers, against the CPU restatement of its published algorithm (oracle/radam_sf_ref.py; the package
itself is absent: parity unpinned for this optimizer), and the on-disk ``model.pth`` layout of main.py:219-226 /
parts/def_for_main.py:63-70 as a save -> load round trip."""
import os

import numpy as np
import pytest
import torch

from oracle import egnn_ref
from oracle.radam_sf_ref import RAdamScheduleFreeRef

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _net(L=2, d=64, seed=0):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    torch.manual_seed(seed)
    return EquivariantGNN(**egnn_ref.net_sizes(L, d)).to(DEV)


@pytest.mark.parametrize("weight_decay", [0.0, 1e-2])
def test_radam_schedulefree_kernel_matches_restatement(lib, weight_decay):
    from diffusion_model_b200.DataPreprocessor import SpectrumCompressor
    from diffusion_model_b200.optim import RAdamScheduleFree
    net = _net()
    comp = SpectrumCompressor(40, [24, 16], 8).to(DEV)  # parameters outside the flat buffer (def_for_main.py:123-127)
    params = list(net.parameters()) + list(comp.parameters())
    ref_p = [p.detach().cpu().clone() for p in params]
    opt = RAdamScheduleFree(params, lr=1e-2, weight_decay=weight_decay)
    ref = RAdamScheduleFreeRef(ref_p, lr=1e-2, weight_decay=weight_decay)
    with pytest.raises(Exception):
        opt.step()  # not in train mode, like the package
    gen = torch.Generator().manual_seed(1)
    for epoch in range(3):
        opt.train(); ref.train()
        for it in range(9):
            for p, q in zip(params, ref_p):
                g = torch.randn(q.shape, generator=gen) * (1.0 + 0.1 * it)
                q.grad = g
                p.grad = g.to(DEV)
            opt.step(); ref.step()
            if epoch == 0 and it < 4:  # rho_t <= 4: the silent phase leaves y and z alone
                assert all(torch.equal(p.detach().cpu(), q0) for p, q0 in zip(params, [r for r in ref_p]))
        for p, q in zip(params, ref_p):
            assert float((p.detach().cpu() - q).abs().max()) <= 2e-6 * max(1.0, float(q.abs().max())), "y after train"
        opt.eval(); ref.eval()
        for p, q in zip(params, ref_p):
            assert float((p.detach().cpu() - q).abs().max()) <= 2e-6 * max(1.0, float(q.abs().max())), "x after eval()"
    g = opt.param_groups[0]
    assert g["k"] == ref.g["k"] == 27 and abs(g["weight_sum"] - ref.g["weight_sum"]) <= 1e-12 * ref.g["weight_sum"]
    assert abs(g["scheduled_lr"] - ref.g["scheduled_lr"]) <= 1e-15 and not g["train_mode"]
    # eval() and train() are inverse maps (x <-> y through z)
    before = [p.detach().clone() for p in params]
    opt.train(); opt.eval()
    assert all(float((a - p.detach()).abs().max()) <= 1e-6 for a, p in zip(before, params))


def test_radam_schedulefree_trains_the_denoiser_and_invalidates_packed_weights(lib):
    """train_epoch / eval_epoch with params['optimizer'] == 'RAdamScheduleFree' (parts/train_per_iretation.py:112-113,
    198-199): the loops switch the optimizer's mode; the kernels must see y while training and x while evaluating."""
    import random
    from diffusion_model_b200.data import Data, DataLoader
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.graph import build_edges_fc
    from diffusion_model_b200.optim import RAdamScheduleFree
    from diffusion_model_b200.train_per_iretation import eval_epoch, train_epoch
    net = _net(2, 64, seed=3)
    net.precision = "bf16x3"
    gen = torch.Generator().manual_seed(5)
    data = []
    for g in range(6):
        n = 16
        x = torch.nn.functional.one_hot(torch.randint(0, 2, (n,), generator=gen), 2).float()
        data.append(Data(x=x, pos=3.0 * torch.rand(n, 3, generator=gen), edge_index=build_edges_fc([n], DEV).cpu()))
    loader = DataLoader(data, batch_size=3)
    dp = E3DiffusionProcess(s=1e-5, power=2.0, num_diffusion_timestep=1000)
    prm = dict(atom_type_size=2, optimizer="RAdamScheduleFree")
    opt = RAdamScheduleFree(net.parameters(), lr=1e-3)
    random.seed(1); torch.manual_seed(1)
    w0 = net.flat_parameters().clone()
    losses = [train_epoch({"egnn": net}, loader, prm, dp, opt) for _ in range(6)]
    assert opt.param_groups[0]["train_mode"] and all(np.isfinite(losses))
    y = net.flat_parameters().clone()
    assert not torch.equal(y, w0)
    random.seed(2); torch.manual_seed(2)
    e1 = eval_epoch({"egnn": net}, loader, prm, dp, opt)
    x_avg = net.flat_parameters().clone()
    assert not opt.param_groups[0]["train_mode"] and not torch.equal(x_avg, y)
    # the same evaluation with the packed images forced stale would differ: re-evaluating reproduces e1 bit for bit
    random.seed(2); torch.manual_seed(2)
    e2 = eval_epoch({"egnn": net}, loader, prm, dp, opt)
    assert e1 == e2
    # and it is the loss AT x, not at y: putting y back by hand (through .data, no version bump) changes the result
    net.flat_parameters().data.copy_(y)
    net.invalidate_packed()
    for layer in net.egcl_list:
        layer.invalidate_packed()
    random.seed(2); torch.manual_seed(2)
    e3 = eval_epoch({"egnn": net}, loader, prm, dp, opt)
    assert e3 != e1


def test_packed_weights_follow_data_updates(lib):
    """parameter writes through ``p.data`` bump no autograd version.  A forward that records a graph always
    re-packs; under no_grad the cache is kept until invalidate_packed() (documented)."""
    from diffusion_model_b200.graph import build_edges_fc
    net = _net(1, 64, seed=4)
    net.precision = "bf16x3"
    ei = build_edges_fc([20], DEV)
    gen = torch.Generator().manual_seed(0)
    h, x = torch.randn(20, 3, generator=gen).to(DEV), torch.randn(20, 3, generator=gen).to(DEV)
    h1, _ = net(ei, h, x)  # grad mode, parameters require grad
    for p in net.parameters():
        p.data.mul_(1.5)  # EMA / weight-swap style update: _version unchanged
    h2, _ = net(ei, h, x)
    assert not torch.equal(h1, h2)
    with torch.no_grad():
        h3, _ = net(ei, h, x)
        assert torch.equal(h3, h2.detach())
        for p in net.parameters():
            p.data.mul_(0.5)
        net.invalidate_packed()
        h4, _ = net(ei, h, x)
    assert not torch.equal(h4, h3)


def test_model_pth_round_trip(lib, tmp_path):
    """main.py:219-226 writes {'egnn': state_dict[, 'spectrum_compressor': ...]} to model.pth; load_model_state
    (parts/def_for_main.py:63-70) reads it back with weights_only=True.  Same keys, same tensors, same outputs."""
    from diffusion_model_b200.DataPreprocessor import SpectrumCompressor
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.graph import build_edges_fc
    net = _net(3, 64, seed=7)
    comp = SpectrumCompressor(200, [150, 100, 50], 32).to(DEV)
    states = {"egnn": net.state_dict(), "spectrum_compressor": comp.state_dict()}
    path = os.path.join(tmp_path, "model.pth")
    torch.save(states, path)
    loaded = torch.load(path, weights_only=True)
    want = [f"egcl_list.{l}.{m}.{i}.{wb}" for l in range(3)
            for m, idx in (("mlp_m", (0, 2)), ("mlp_x", (0, 2, 4)), ("mlp_h", (0, 2)), ("attention", (0,)))
            for i in idx for wb in ("weight", "bias")]
    assert list(loaded["egnn"].keys()) == want  # the reference module's named_parameters() order
    torch.manual_seed(99)
    net2 = EquivariantGNN(**egnn_ref.net_sizes(3, 64)).to(DEV)
    comp2 = SpectrumCompressor(200, [150, 100, 50], 32).to(DEV)
    net2.load_state_dict(loaded["egnn"])
    comp2.load_state_dict(loaded["spectrum_compressor"])
    assert all(torch.equal(a, b) for a, b in zip(net.state_dict().values(), net2.state_dict().values()))
    ei = build_edges_fc([24, 24], DEV)
    gen = torch.Generator().manual_seed(0)
    h, x = torch.randn(48, 3, generator=gen).to(DEV), torch.randn(48, 3, generator=gen).to(DEV)
    for precision in ("fp32", "bf16x3"):
        net.precision = net2.precision = precision
        with torch.no_grad():
            a, b = net(ei, h, x), net2(ei, h, x)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    # a state_dict saved by the reference module (CPU tensors, same keys) loads too
    sd_cpu = egnn_ref.init_state_dict(**egnn_ref.net_sizes(3, 64))
    torch.save({"egnn": sd_cpu}, path)
    net2.load_state_dict(torch.load(path, weights_only=True)["egnn"])
    with torch.no_grad():
        hr, xr = egnn_ref.egnn_forward(sd_cpu, 3, ei.cpu(), h.cpu(), x.cpu())
        net2.precision = "fp32"
        ho, xo = net2(ei, h, x)
    assert float((ho.cpu() - hr).abs().max() / hr.abs().max()) <= 3e-4
