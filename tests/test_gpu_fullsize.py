"""Full-size checks (BASELINE.json cfg2 and cfg4 as bench.py builds them) through size-independent properties -- the
oracle cannot run these sizes in seconds, so nothing here compares against it:
  * reruns of forward + backward are bit-identical (fixed-order reductions, no atomics);
  * E(3) equivariance of the denoiser (EquivariantGraphNeuralNetwork.py:55-71): h' invariant, x' rotates and shifts with x;
  * the tcgen05 bf16x3 mode agrees with the CUDA-core fp32 mode (which the golden-size tests pin to the oracle): outputs
    within 1e-3 over all L layers, every gradient tensor within 1e-3 or within what a 1e-4 relative perturbation of the
    inputs does to that tensor (backward stability; some tensors are ill-conditioned at random init)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _workload(name):
    import bench
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    from diffusion_model_b200.graph import build_edges_cutoff
    cfg = dict(bench.CFG)
    if name == "cfg4":
        cfg.update(bench.CFG4)
    G, n = cfg["graphs"], cfg["atoms"]
    pos, onehot, box = bench.synth_batch(G, n, 77)
    node_ptr = (torch.arange(G + 1) * n).to(torch.int32).to(DEV)
    pos_d = pos.to(DEV)
    ei = build_edges_cutoff(pos_d, cfg["r_c"], node_ptr=node_ptr, box=box.to(DEV))
    gen = torch.Generator().manual_seed(5)
    tcol = torch.rand(G, generator=gen).repeat_interleave(n).unsqueeze(1)  # t/T, one value per graph
    h = torch.cat([onehot + 0.3 * torch.randn(onehot.shape, generator=gen), tcol], dim=1).to(DEV)
    torch.manual_seed(11)
    net = EquivariantGNN(**bench.net_sizes(cfg["L"], cfg["d"], cfg["H"])).to(DEV)
    return net, ei, h, pos_d, node_ptr


def _fwd_bwd(net, ei, h, x, node_ptr, precision):
    net.precision = precision
    for p in net.parameters():
        p.grad = None
    hc, xc = h.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ho, xo = net(ei, hc, xc, node_ptr=node_ptr)
    # a coherent objective (like the eps-prediction loss): with random-sign output weights every weight gradient is a
    # cancelling sum over 4e5 edges and max-abs relative errors measure the cancellation, not the kernels
    (ho.square().sum() + (xo - x).square().sum()).backward()
    grads = {k: p.grad.clone() for k, p in net.named_parameters()}
    grads["h_in"], grads["x_in"] = hc.grad.clone(), xc.grad.clone()
    return ho.detach(), xo.detach(), grads


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("name", ["cfg2", "cfg4"])
def test_fullsize_rerun_bit_identical_and_modes_agree(lib, name):
    net, ei, h, x, node_ptr = _workload(name)
    h1, x1, g1 = _fwd_bwd(net, ei, h, x, node_ptr, "bf16x3")
    h2, x2, g2 = _fwd_bwd(net, ei, h, x, node_ptr, "bf16x3")
    assert torch.equal(h1, h2) and torch.equal(x1, x2)
    assert all(torch.equal(g1[k], g2[k]) for k in g1)
    assert all(bool(torch.isfinite(v).all()) for v in g1.values())
    hf, xf, gf = _fwd_bwd(net, ei, h, x, node_ptr, "fp32")
    eh, ex = _rel(h1, hf), _rel(x1 - x, xf - x)
    errs = {k: _rel(g1[k], gf[k]) for k in g1}
    # conditioning of each gradient tensor: its response (fp32 mode) to a 1e-6 relative perturbation of the inputs.  At
    # random init some mid-stack tensors amplify 200x (scripts/dbg_grad_condition.py: the bf16x3 difference tracks this
    # response at a constant ratio of ~34, i.e. it equals an input perturbation of 3.4e-5), so the bound is "no worse
    # than a 1e-4 relative input perturbation", with the golden-size parity bar of 1e-3 as the floor
    gen = torch.Generator().manual_seed(1)
    hp = h * (1 + 1e-6 * torch.randn(h.shape, generator=gen).to(DEV))
    xp = x * (1 + 1e-6 * torch.randn(x.shape, generator=gen).to(DEV))
    _, _, gp = _fwd_bwd(net, ei, hp, xp, node_ptr, "fp32")
    resp = {k: _rel(gp[k], gf[k]) for k in g1}
    worst = max(errs, key=errs.get)
    print(f"full-size {name}: E={ei.shape[1]} bf16x3 vs fp32 mode: h {eh:.2e} dx {ex:.2e} worst grad {worst} {errs[worst]:.2e} "
          f"(response to 1e-6: {resp[worst]:.2e})")
    assert eh <= 1e-3 and ex <= 1e-3, (eh, ex)  # 1e-4 per layer (tests/test_gpu_tc.py), L layers deep
    bad = {k: (errs[k], resp[k]) for k in errs if errs[k] > max(1e-3, 100.0 * resp[k])}
    assert not bad, bad


@pytest.mark.parametrize("name", ["cfg2", "cfg4"])
def test_fullsize_equivariance(lib, name):
    net, ei, h, x, node_ptr = _workload(name)
    net.precision = "bf16x3"
    rng = np.random.default_rng(9)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    R = torch.from_numpy(q).float().to(DEV)
    shift = torch.tensor([0.7, -1.3, 2.1], device=DEV)
    with torch.no_grad():
        h0, x0 = net(ei, h, x, node_ptr=node_ptr)
        x_rot = x @ R.t() + shift
        h1, x1 = net(ei, h, x_rot, node_ptr=node_ptr)
    eh = _rel(h1, h0)
    ex = float(((x1 - x_rot) - (x0 - x) @ R.t()).abs().max() / (x0 - x).abs().max().clamp_min(1e-30))
    print(f"full-size {name}: equivariance h {eh:.2e} dx {ex:.2e}")
    # the rotated coordinates are re-rounded to fp32 (cell side up to 28.5 A) before ||x_i - x_j||^2 is formed
    assert eh <= 2e-3 and ex <= 2e-3, (eh, ex)
