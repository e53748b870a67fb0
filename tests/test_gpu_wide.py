"""Tile-GEMM tcgen05 forward (csrc/tc_wide.cu) of the edge MLPs for the network shapes the fused kernels do not cover
(EquivariantGraphNeuralNetwork.py:55-65 with the widths of parameters.yaml / main.py:102-121, and ragged widths).
Against the CPU oracle on the oracle's own layer inputs: bf16x3 1e-4 per layer (the fp32 bar), bf16 the tensor-core bound of
tests/test_gpu_tc.py; gradients (fp32 backward on the forward's saved sums) 1e-3."""
import pytest
import torch

from oracle import egnn_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def relerr(a, b):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _sizes(L, H, Mh, Ms, Xh, Hh):
    return dict(L=L, m_input=2 * H + 1, m_hidden=Mh, m_output=Ms, x_input=2 * H + 1, x_hidden=Xh, x_output=1,
                h_input=H + Ms, h_hidden=Hh, h_output=H)


def _graph(n_per_graph, H, seed, drop_node=None):
    """fully connected graphs; drop_node loses all its incoming edges (a receiver with an empty edge range)"""
    from diffusion_model_b200.graph import build_edges_fc
    gen = torch.Generator().manual_seed(seed)
    N = sum(n_per_graph)
    h0, x0 = torch.randn(N, H, generator=gen), 2.0 * torch.randn(N, 3, generator=gen)
    ei = build_edges_fc(n_per_graph, DEV).cpu()
    if drop_node is not None:
        ei = ei[:, ei[0] != drop_node]
    return h0, x0, ei


SHAPES = {
    # name: (H, Mh, Ms, Xh, Hh, graphs)
    "yaml": (36, 1024, 256, 1024, 1024, [40, 33]),        # 2 682 edges: 21 edge tiles, the last one ragged
    "ragged": (5, 200, 72, 328, 96, [17, 9, 30]),         # every width padded: 11 -> 64, 200 -> 256, 328 -> 512, 72 -> 256
    "narrow": (9, 48, 40, 24, 32, [12, 7]),               # a single partial tile everywhere
    "k128": (63, 320, 256, 256, 128, [20]),               # K0 = 127: two full K slices of the gather image
}


@pytest.mark.parametrize("shape", list(SHAPES))
@pytest.mark.parametrize("precision", ["bf16x3", "bf16"])
def test_wide_forward_layers_match_oracle(lib, shape, precision):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    H, Mh, Ms, Xh, Hh, n = SHAPES[shape]
    L = 2
    sizes = _sizes(L, H, Mh, Ms, Xh, Hh)
    torch.manual_seed(5)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), L)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    h0, x0, ei = _graph(n, H, 6, drop_node=3)
    with torch.no_grad():
        _, _, layers = egnn_ref.egnn_forward(sd, L, ei, h0, x0, return_layers=True)
    bound = 1e-4 if precision == "bf16x3" else 1.5e-1
    h, x, worst = h0, x0, 0.0
    with torch.no_grad():
        for l, (hl, xl) in enumerate(layers):
            layer = net.egcl_list[l]
            layer.precision = precision
            ho, xo = layer(ei.to(DEV), h.to(DEV), x.to(DEV))
            eh, ex = relerr(ho, hl), relerr(xo, xl)
            assert eh <= bound and ex <= bound, (shape, l, eh, ex)
            worst = max(worst, eh, ex)
            h, x = hl, xl
    print(f"wide forward {shape} {precision}: worst per-layer error {worst:.2e}")


def test_wide_network_gradients_match_oracle(lib):
    """bf16x3 forward through the tile-GEMM path, backward through the fp32 kernels, whole network (3 layers, ragged widths)"""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    H, Mh, Ms, Xh, Hh, n = SHAPES["ragged"]
    L = 3
    sizes = _sizes(L, H, Mh, Ms, Xh, Hh)
    torch.manual_seed(8)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), L)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    net.precision = "bf16x3"
    h0, x0, ei = _graph(n, H, 9)
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    hc, xc = h0.clone().requires_grad_(True), x0.clone().requires_grad_(True)
    hr, xr = egnn_ref.egnn_forward(sdg, L, ei, hc, xc)
    (hr.square().sum() + (xr - xc).square().sum()).backward()
    hd, xd = h0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
    ho, xo = net(ei.to(DEV), hd, xd)
    assert relerr(ho, hr) <= 3e-4 and relerr(xo, xr) <= 3e-4, (relerr(ho, hr), relerr(xo, xr))
    (ho.square().sum() + (xo - xd).square().sum()).backward()
    errs = {k: relerr(p.grad, sdg[k].grad) for k, p in net.named_parameters()}
    errs["h_in"], errs["x_in"] = relerr(hd.grad, hc.grad), relerr(xd.grad, xc.grad)
    worst = max(errs, key=errs.get)
    print(f"wide path gradients: worst {worst} {errs[worst]:.2e}")
    assert errs[worst] <= 1e-3, sorted(errs.items(), key=lambda kv: -kv[1])[:5]


def test_wide_rerun_is_bitwise_identical(lib):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    H, Mh, Ms, Xh, Hh, n = SHAPES["yaml"]
    sizes = _sizes(1, H, Mh, Ms, Xh, Hh)
    torch.manual_seed(10)
    net = EquivariantGNN(**sizes).to(DEV)
    net.precision = "bf16x3"
    h0, x0, ei = _graph(n, H, 11)
    with torch.no_grad():
        a = net(ei.to(DEV), h0.to(DEV), x0.to(DEV))
        b = net(ei.to(DEV), h0.to(DEV), x0.to(DEV))
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_outside_the_wide_limits_is_refused(lib):
    """K0 = 141 > 128 has neither a fused kernel nor a tile-GEMM layout: the library must refuse bf16x3, not fall back"""
    from diffusion_model_b200 import _lib
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    H = 70
    net = EquivariantGNN(**_sizes(1, H, 64, 64, 64, 64)).to(DEV)
    assert net.resolved_precision() == "fp32"
    net.precision = "bf16x3"
    h0, x0, ei = _graph([6], H, 12)
    with pytest.raises(_lib.EgnnError):
        net(ei.to(DEV), h0.to(DEV), x0.to(DEV))


@pytest.mark.parametrize("d,Hh", [(128, 192), (64, 40), (256, 256)])
def test_fused_edge_kernels_with_tile_gemm_node_mlp(lib, d, Hh):
    """d in {64, 128} with a node-MLP width the fused tcgen05 node kernels do not cover (Hh != d), and d = 256 (no fused node
    kernel at all): the edge phase runs the fused kernels, whose per-receiver sums are cut at 32-edge segment boundaries, and
    the node MLP the tile GEMMs (forward: wide_combine + [h | M] image; backward: egnn_wide_node_backward)."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    H, L = 3, 2
    sizes = _sizes(L, H, d, d, d, Hh)
    torch.manual_seed(21)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), L)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    assert net.resolved_precision() == "bf16x3"
    h0, x0, ei = _graph([40, 51, 9], H, 22, drop_node=5)   # receivers with 39 / 50 / 8 edges: ranges straddle the 32-edge segments
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    hc, xc = h0.clone().requires_grad_(True), x0.clone().requires_grad_(True)
    hr, xr = egnn_ref.egnn_forward(sdg, L, ei, hc, xc)
    (hr.square().sum() + (xr - xc).square().sum()).backward()
    hd, xd = h0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
    ho, xo = net(ei.to(DEV), hd, xd)
    assert relerr(ho, hr) <= 2e-4 and relerr(xo, xr) <= 2e-4, (relerr(ho, hr), relerr(xo, xr))
    (ho.square().sum() + (xo - xd).square().sum()).backward()
    errs = {k: relerr(p.grad, sdg[k].grad) for k, p in net.named_parameters()}
    errs["h_in"], errs["x_in"] = relerr(hd.grad, hc.grad), relerr(xd.grad, xc.grad)
    worst = max(errs, key=errs.get)
    print(f"fused edge + tile-GEMM node, d={d} Hh={Hh}: fwd {relerr(ho, hr):.1e}, worst gradient {worst} {errs[worst]:.2e}")
    assert errs[worst] <= 1e-3, sorted(errs.items(), key=lambda kv: -kv[1])[:5]


def test_tile_gemm_edges_with_cuda_core_node_mlp(lib):
    """h_hidden = 1100 is past the tile-GEMM limit of the node MLP (1024): the edge MLPs still run on the tile GEMMs (forward
    and backward), the node MLP on the CUDA-core kernels, which read the per-receiver sums and phi_e the edge phase leaves."""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    H, L = 7, 2
    sizes = _sizes(L, H, 300, 96, 260, 1100)
    torch.manual_seed(31)
    sd = egnn_ref.tame_state_dict(egnn_ref.init_state_dict(**sizes), L)
    net = EquivariantGNN(**sizes)
    net.load_state_dict(sd)
    net = net.to(DEV)
    assert net.resolved_precision() == "bf16x3"
    h0, x0, ei = _graph([23, 14], H, 32, drop_node=2)
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    hc, xc = h0.clone().requires_grad_(True), x0.clone().requires_grad_(True)
    hr, xr = egnn_ref.egnn_forward(sdg, L, ei, hc, xc)
    (hr.square().sum() + (xr - xc).square().sum()).backward()
    hd, xd = h0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
    ho, xo = net(ei.to(DEV), hd, xd)
    assert relerr(ho, hr) <= 2e-4 and relerr(xo, xr) <= 2e-4, (relerr(ho, hr), relerr(xo, xr))
    (ho.square().sum() + (xo - xd).square().sum()).backward()
    errs = {k: relerr(p.grad, sdg[k].grad) for k, p in net.named_parameters()}
    errs["h_in"], errs["x_in"] = relerr(hd.grad, hc.grad), relerr(xd.grad, xc.grad)
    worst = max(errs, key=errs.get)
    print(f"tile-GEMM edges + CUDA-core node MLP: fwd {relerr(ho, hr):.1e}, worst gradient {worst} {errs[worst]:.2e}")
    assert errs[worst] <= 1e-3, sorted(errs.items(), key=lambda kv: -kv[1])[:5]
