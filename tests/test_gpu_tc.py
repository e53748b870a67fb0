"""tcgen05 path: one-tile GEMM self-test (pins descriptor conventions on hardware) and, further down, parity of the
tensor-core edge kernels against the fp32 reference results."""
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _selftest(K, N, npass, swap):
    from diffusion_model_b200._lib import call, ptr, stream_ptr
    gen = torch.Generator().manual_seed(K * 1000 + N)
    A = torch.randn(128, K, generator=gen).to(DEV)
    B = torch.randn(N, K, generator=gen).to(DEV)
    D = torch.full((128, N), float("nan"), device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    call("egnn_tc_selftest", ptr(A), ptr(B), ptr(D), K, N, npass, swap, ptr(status), stream_ptr())
    torch.cuda.synchronize()
    ref = A.double() @ B.double().t()
    err = float((D.double() - ref).abs().max() / ref.abs().max())
    return err, int(status.item())


def test_tc_selftest_descriptor_convention(lib):
    out = {}
    os.makedirs("gpurun_out", exist_ok=True)
    for swap in (0, 1):
        for K, N in ((16, 16), (16, 256), (64, 64), (128, 128)):
            for npass in (1, 3):
                err, st = _selftest(K, N, npass, swap)
                out[f"K{K}_N{N}_p{npass}_swap{swap}"] = (err, st)
                json.dump(out, open("gpurun_out/tc_selftest.json", "w"), indent=1)
    # MN-major operands (backward kernels view an edge tile [e][n] as the transposed operand): probe both conventions
    for flags in (2, 2 | 8, 4, 4 | 8, 6, 6 | 8):
        for K, N in ((16, 16), (64, 64), (128, 128), (128, 16)):
            err, st = _selftest(K, N, 3, flags)
            out[f"K{K}_N{N}_p3_flags{flags}"] = (err, st)
            json.dump(out, open("gpurun_out/tc_selftest.json", "w"), indent=1)
    for K, N in ((16, 16), (16, 256), (64, 64), (128, 128)):
        e1, s1 = out[f"K{K}_N{N}_p1_swap0"]
        e3, s3 = out[f"K{K}_N{N}_p3_swap0"]
        assert s1 == 0 and s3 == 0, out
        assert e1 < 2e-2, out   # single bf16 product
        assert e3 < 2e-5, out   # hi/lo split, 3 products


# ---------------------------------------------------------------- tensor-core edge kernel vs the reference outputs
import numpy as np  # noqa: E402

from oracle import egnn_ref  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(GOLD, name + ".npz"), allow_pickle=False)


def _net_from(z):
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    sizes = {str(k): int(v) for k, v in zip(z["size_keys"], z["sizes"])}
    net = EquivariantGNN(**sizes)
    net.load_state_dict({k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w:")})
    return net.to(DEV), sizes


def _relerr(a, b):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg4"])  # cfg4: d = 256, the streamed-W2 kernel (tc_edge256.cu)
@pytest.mark.parametrize("precision,tol", [("bf16x3", 1e-4), ("bf16", 2e-2)])
def test_tc_forward_matches_reference_layers(lib, name, precision, tol):
    z = _load("egnn_" + name)
    net, sizes = _net_from(z)
    h, x = torch.from_numpy(z["h"]).to(DEV), torch.from_numpy(z["x"]).to(DEV)
    ei, batch = torch.from_numpy(z["edge_index"]).to(DEV), torch.from_numpy(z["batch"]).to(DEV)
    worst = 0.0
    with torch.no_grad():
        for l in range(sizes["L"]):
            layer = net.egcl_list[l]
            layer.precision = precision
            hh, xx = layer(ei, torch.from_numpy(z[f"h_l{l - 1}"]).to(DEV) if l else h,
                           torch.from_numpy(z[f"x_l{l - 1}"]).to(DEV) if l else x)
            eh, ex = _relerr(hh, z[f"h_l{l}"]), _relerr(xx, z[f"x_l{l}"])
            worst = max(worst, eh, ex)
            assert eh <= tol and ex <= tol, (name, precision, l, eh, ex)
        net.precision = precision
        hg, xg = net(ei, h, x, batch=batch, norm_scope="graph")
        h2, x2 = net(ei, h, x, batch=batch, norm_scope="graph")
        assert torch.equal(hg, h2) and torch.equal(xg, x2)  # deterministic
        if precision == "bf16x3":
            assert _relerr(hg, z["h_graph"]) <= 5e-4 and _relerr(xg, z["x_graph"]) <= 5e-4  # L layers compounded
    print(f"tc forward {name} {precision}: worst per-layer relerr {worst:.2e}")


def test_tc_forward_edge_cases_and_unsupported(lib):
    from diffusion_model_b200 import EgnnError
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    torch.manual_seed(0)
    net = EquivariantGNN(**egnn_ref.net_sizes(2, 64)).to(DEV)
    net.precision = "bf16x3"
    sd = {k: v.detach().cpu() for k, v in net.state_dict().items()}
    gen = torch.Generator().manual_seed(1)
    n = 320
    recv = np.concatenate([np.full(300, 2), np.full(5, 7), np.full(200, 9)]).astype(np.int64)
    send = np.concatenate([np.arange(10, 310), np.arange(0, 5), np.arange(100, 300)]).astype(np.int64)
    ei = torch.from_numpy(np.stack([recv, send]))
    h, x = torch.randn(n, 3, generator=gen), torch.randn(n, 3, generator=gen)
    with torch.no_grad():
        ho, xo = net(ei.to(DEV), h.to(DEV), x.to(DEV))
        hr, xr = egnn_ref.egnn_forward(sd, 2, ei, h, x)
    assert _relerr(ho, hr) <= 2e-4 and _relerr(xo, xr) <= 2e-4
    # d = 256 on the same ragged graph (one receiver cut by several tiles, a 5-edge receiver, isolated nodes), both modes
    torch.manual_seed(3)
    wide = EquivariantGNN(**egnn_ref.net_sizes(2, 256)).to(DEV)
    sdw = {k: v.detach().cpu() for k, v in wide.state_dict().items()}
    with torch.no_grad():
        hr, xr = egnn_ref.egnn_forward(sdw, 2, ei, h, x)
        for prec, tol in (("bf16x3", 2e-4), ("bf16", 3e-2)):
            wide.precision = prec
            ho, xo = wide(ei.to(DEV), h.to(DEV), x.to(DEV))
            assert _relerr(ho, hr) <= tol and _relerr(xo, xr) <= tol, (prec, _relerr(ho, hr), _relerr(xo, xr))
            ho2, xo2 = wide(ei.to(DEV), h.to(DEV), x.to(DEV))
            assert torch.equal(ho, ho2) and torch.equal(xo, xo2)
    # d = 96 has no fused kernel: the tile-GEMM path (csrc/tc_wide.cu) pads it to 256 columns
    odd = EquivariantGNN(**egnn_ref.net_sizes(1, 96)).to(DEV)
    sdo = {k: v.detach().cpu() for k, v in odd.state_dict().items()}
    odd.precision = "bf16x3"
    with torch.no_grad():
        hr, xr = egnn_ref.egnn_forward(sdo, 1, ei, h, x)
        ho, xo = odd(ei.to(DEV), h.to(DEV), x.to(DEV))
    assert _relerr(ho, hr) <= 1e-4 and _relerr(xo, xr) <= 1e-4, (_relerr(ho, hr), _relerr(xo, xr))
    # K0 = 2H+1 = 141 > 128 is outside both tensor-core paths: must refuse, not fall back
    H = 70
    far = EquivariantGNN(**egnn_ref.net_sizes(1, 64, H=H)).to(DEV)
    far.precision = "bf16"
    with pytest.raises(EgnnError):
        far(ei.to(DEV), torch.randn(h.shape[0], H, device=DEV), x.to(DEV))


# ---------------------------------------------------------------- tensor-core edge BACKWARD vs the reference autograd
def _grad_errors(name, precision, scope="call"):
    z = _load("egnn_" + name)
    net, sizes = _net_from(z)
    net.precision = precision
    sd = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w:")}
    h0, x0 = torch.from_numpy(z["h"]), torch.from_numpy(z["x"])
    ei = torch.from_numpy(z["edge_index"])
    th, tx = torch.from_numpy(z["target_h"]), torch.from_numpy(z["target_x"])
    h, x = h0.to(DEV).requires_grad_(True), x0.to(DEV).requires_grad_(True)
    ho, xo = net(ei.to(DEV), h, x)
    loss = ((ho - th.to(DEV)) ** 2).sum() + ((xo - x - tx.to(DEV)) ** 2).sum()
    loss.backward()
    errs = {k: _relerr(p.grad, z["g:" + k]) for k, p in net.named_parameters()}
    hc, xc = h0.clone().requires_grad_(True), x0.clone().requires_grad_(True)
    hr, xr = egnn_ref.egnn_forward(sd, sizes["L"], ei, hc, xc)
    (((hr - th) ** 2).sum() + ((xr - xc - tx) ** 2).sum()).backward()
    errs["h_in"], errs["x_in"] = _relerr(h.grad, hc.grad), _relerr(x.grad, xc.grad)
    errs["loss"] = abs(loss.item() - float(z["loss"])) / abs(float(z["loss"]))
    grads = [p.grad.clone() for p in net.parameters()] + [h.grad.clone(), x.grad.clone()]
    return errs, grads


@pytest.mark.parametrize("name", ["cfg1", "cfg2"])
# bf16 (ONE tcgen05 product per MMA, no lo halves) is the stated looser mode: its gradient bound is what was measured on
# B200 -- worst tensor 3.1e-2 (cfg1), 9.9e-2 (cfg2), 9.4e-2 (d = 256) -- with 1.5x headroom, not a round number
@pytest.mark.parametrize("precision,tol", [("bf16x3", 1e-3), ("bf16", 1.5e-1)])
def test_tc_backward_matches_reference_grads(lib, name, precision, tol):
    errs, g1 = _grad_errors(name, precision)
    worst = max(errs, key=errs.get)
    print(f"tc backward {name} {precision}: worst {worst} {errs[worst]:.2e}")
    assert errs[worst] <= tol, errs
    _, g2 = _grad_errors(name, precision)
    assert all(torch.equal(a, b) for a, b in zip(g1, g2))  # deterministic (no atomics)


@pytest.mark.parametrize("precision,tol", [("bf16x3", 1e-3), ("bf16", 1.5e-1)])
def test_tc_backward_d256(lib, precision, tol):
    """d = 256 (cfg4 net): the two-kernel tcgen05 backward of tc_edge256_bwd.cu (tile images + split-K wgrad GEMM)"""
    errs, g1 = _grad_errors("cfg4", precision)
    worst = max(errs, key=errs.get)
    top = sorted(errs.items(), key=lambda kv: -kv[1])[:8]
    print(f"tc backward d256 {precision}: worst {worst} {errs[worst]:.2e}", [(k, f"{v:.1e}") for k, v in top])
    assert errs[worst] <= tol, top
    _, g2 = _grad_errors("cfg4", precision)
    assert all(torch.equal(a, b) for a, b in zip(g1, g2))


@pytest.mark.parametrize("d", [64, 256])  # 256: the two-kernel backward (tile images + wgrad GEMM), fewer tiles than CTAs
def test_tc_backward_edge_cases(lib, d):
    """a receiver cut by several tiles, isolated nodes, one short tile, no edges"""
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    torch.manual_seed(0)
    net = EquivariantGNN(**egnn_ref.net_sizes(2, d)).to(DEV)
    net.precision = "bf16x3"
    sd = {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    gen = torch.Generator().manual_seed(1)
    n = 320
    recv = np.concatenate([np.full(300, 2), np.full(5, 7), np.full(200, 9)]).astype(np.int64)
    send = np.concatenate([np.arange(10, 310), np.arange(0, 5), np.arange(100, 300)]).astype(np.int64)
    for ei in (torch.from_numpy(np.stack([recv, send])), torch.from_numpy(np.stack([recv[:40], send[:40]])),
               torch.zeros(2, 0, dtype=torch.int64)):
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
            scale = max(float(ref.abs().max()), 1e-2)  # scalar bias gradients are sums with heavy cancellation
            assert float((p.grad.cpu() - ref).abs().max()) <= 1e-3 * scale, (k, ei.shape)
        assert _relerr(h.grad, hc.grad) <= 1e-3 and _relerr(x.grad, xc.grad) <= 1e-3
