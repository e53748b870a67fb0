"""Full-size parity: the workloads bench.py times (BASELINE.json cfg2: 64 x 192-atom periodic cutoff graphs, L=6, d=128;
cfg4: 1536-atom cells, L=9, d=256 -- two graphs of it, so that every CTA of the d=256 kernels walks several tiles)
against the CPU oracle on the same seeded inputs:

  * every EGCL layer on the oracle's own layer input: h', x' within 1e-4 (max |err| / max |ref|), both precisions;
  * the eps-prediction loss of the minibatch (parts/train_per_iretation.py:130-169) with injected t / noise: 1e-5;
  * every parameter gradient: 1e-3 of the tensor's largest entry.

cfg2 uses bench.py's own seeds (batch 2025, weights 2024).  cfg4 is checked twice: with the tamed weights of the sampling
fixtures (bounded activations: the flat 1e-3 bar applies), and with bench.py's default-init weights, where activations reach
1e3 by layer 5 (no residual on h, EquivariantGraphNeuralNetwork.py:69; the author's own note that deep stacks give NaN,
train.py:101).  There the yardstick is the reference arithmetic's OWN fp32 round-off on that tensor -- the fp32 oracle against
the float64 oracle -- and a gradient may exceed 1e-3 only by a stated multiple of it.

Also kept: bit-identical reruns (no atomics) and E(3) equivariance at the same sizes.
"""
import time

import numpy as np
import pytest
import torch

from oracle import diffusion_ref as D
from oracle import egnn_ref, graphs_ref as GR, loops_ref as LP

pytestmark = pytest.mark.gpu
DEV = "cuda"
T, S_NOISE, POWER = 1000, 1e-5, 2.0
_CACHE = {}


def _rel(a, b):
    a, b = torch.as_tensor(a).detach().cpu().double(), torch.as_tensor(b).detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _cfg(name):
    import bench
    cfg = dict(bench.CFG)
    if name.startswith("cfg4"):
        cfg.update(bench.CFG4)
        cfg["graphs"] = 2
    return cfg


def _oracle(name, need64=False):
    """Host tensors of the workload, injected draws, and the oracle's loss / per-layer states / gradients (cached)."""
    if name in _CACHE:
        return _CACHE[name]
    import bench
    cfg = _cfg(name)
    G, n, L = cfg["graphs"], cfg["atoms"], cfg["L"]
    pos, onehot, box = bench.synth_batch(G, n, cfg["seed"] + 1)  # rank 0's batch in bench.py
    node_ptr = torch.arange(G + 1) * n
    ei = torch.from_numpy(GR.cutoff_edges_batch(pos.numpy(), node_ptr.numpy(), cfg["r_c"], box.numpy()))
    torch.manual_seed(cfg["seed"])
    sd = egnn_ref.init_state_dict(**bench.net_sizes(L, cfg["d"], cfg["H"]))  # == EquivariantGNN(...) under the same seed
    if name.endswith("tamed"):
        sd = egnn_ref.tame_state_dict(sd, L)
    batch = torch.arange(G).repeat_interleave(n)
    gen = torch.Generator().manual_seed(3)
    t = torch.randint(1, T + 1, (G,), generator=gen)
    noise_pos, noise_h = torch.randn(G * n, 3, generator=gen), torch.randn(G * n, 2, generator=gen)
    sch = D.Schedule(S_NOISE, POWER, T)
    t0 = time.time()
    out = dict(cfg=cfg, pos=pos, onehot=onehot, box=box, ei=ei, sd=sd, batch=batch, t=t, noise_pos=noise_pos,
               noise_h=noise_h, node_ptr=node_ptr)
    for dt in (torch.float32, torch.float64) if need64 else (torch.float32,):
        p = {k: v.to(dt).clone().requires_grad_(True) for k, v in sd.items()}
        loss, aux = LP.train_loss(lambda e, h, x: egnn_ref.egnn_forward(p, L, e, h, x, checkpoint=True), sch, pos.to(dt),
                                  onehot.to(dt), ei, batch, t, noise_pos.to(dt), noise_h.to(dt))
        loss.backward()
        tag = "32" if dt == torch.float32 else "64"
        out["loss" + tag] = float(loss.detach())
        out["grads" + tag] = {k: v.grad.detach().clone() for k, v in p.items()}
        if dt == torch.float32:
            with torch.no_grad():
                _, _, layers = egnn_ref.egnn_forward(sd, L, ei, aux["h_in"].detach(), aux["pos_t"].detach(), return_layers=True)
            out["h_in"], out["pos_t"], out["layers"] = aux["h_in"].detach(), aux["pos_t"].detach(), layers
    out["oracle_s"] = time.time() - t0
    _CACHE[name] = out
    return out


def _net(o, precision):
    import bench
    from diffusion_model_b200.EquivariantGraphNeuralNetwork import EquivariantGNN
    cfg = o["cfg"]
    net = EquivariantGNN(**bench.net_sizes(cfg["L"], cfg["d"], cfg["H"]))
    net.load_state_dict(o["sd"])
    net = net.to(DEV)
    net.precision = precision
    return net


def _gpu_edges(o):
    from diffusion_model_b200.graph import build_edges_cutoff
    cfg = o["cfg"]
    ei = build_edges_cutoff(o["pos"].to(DEV), cfg["r_c"], node_ptr=o["node_ptr"].to(torch.int32).to(DEV), box=o["box"].to(DEV))
    assert torch.equal(ei.cpu(), o["ei"]), "edge_index of the benchmarked batch is not bit-exact"
    return ei


def _layer_errors(o, net, ei):
    errs = []
    h, x = o["h_in"], o["pos_t"]
    np_ = o["node_ptr"].to(torch.int32).to(DEV)
    with torch.no_grad():
        for l, (hr, xr) in enumerate(o["layers"]):
            layer = net.egcl_list[l]
            layer.precision = net.precision
            ho, xo = layer(ei, h.to(DEV), x.to(DEV), node_ptr=np_)
            errs.append((_rel(ho, hr), _rel(xo, xr), _rel(xo.cpu() - x, xr - x)))
            h, x = hr, xr  # the next layer is fed the ORACLE's state: per-layer error, not compounded
    return errs


def _loss_and_grads(o, net, ei):
    from diffusion_model_b200.data import Data
    from diffusion_model_b200.diffusion_x_h import E3DiffusionProcess
    from diffusion_model_b200.train_per_iretation import batch_loss
    dp = E3DiffusionProcess(s=S_NOISE, power=POWER, num_diffusion_timestep=T)
    bd = Data(x=o["onehot"].to(DEV), pos=o["pos"].to(DEV), edge_index=ei)
    bd.node_ptr = o["node_ptr"].to(torch.int32).to(DEV)
    net.zero_grad(set_to_none=True)
    loss, G, _ = batch_loss(net, bd, dp, dict(atom_type_size=2), t=o["t"].tolist(), noise_pos=o["noise_pos"].to(DEV),
                            noise_h=o["noise_h"].to(DEV))
    loss.backward()
    assert G == o["cfg"]["graphs"]
    return float(loss), {k: p.grad.detach().cpu().clone() for k, p in net.named_parameters()}


@pytest.mark.parametrize("precision", ["bf16x3", "fp32"])
@pytest.mark.parametrize("name", ["cfg2", "cfg4_tamed"])
def test_fullsize_matches_oracle(lib, name, precision):
    o = _oracle(name)
    net = _net(o, precision)
    ei = _gpu_edges(o)
    lerr = _layer_errors(o, net, ei)
    loss, grads = _loss_and_grads(o, net, ei)
    eloss = abs(loss - o["loss32"]) / abs(o["loss32"])
    gerr = {k: _rel(grads[k], o["grads32"][k]) for k in grads}
    worst = max(gerr, key=gerr.get)
    print(f"full-size {name} {precision}: E={ei.shape[1]} L={o['cfg']['L']} oracle {o['oracle_s']:.1f}s | per-layer max h "
          f"{max(e[0] for e in lerr):.2e} x {max(e[1] for e in lerr):.2e} dx {max(e[2] for e in lerr):.2e} | loss {eloss:.2e} "
          f"| worst grad {worst} {gerr[worst]:.2e}")
    for l, (eh, ex, edx) in enumerate(lerr):
        assert eh <= 1e-4 and ex <= 1e-4, (l, eh, ex)
        assert edx <= 1e-3, (l, edx)  # the coordinate increment on its own scale
    assert eloss <= 1e-5, (loss, o["loss32"])
    assert gerr[worst] <= 1e-3, sorted(gerr.items(), key=lambda kv: -kv[1])[:6]


@pytest.mark.parametrize("precision,mult", [("bf16x3", 100.0), ("fp32", 10.0)])
def test_fullsize_cfg4_default_init_within_reference_roundoff(lib, precision, mult):
    """bench.py's own cfg4 weights (default init, activations up to 1e3): per-layer outputs keep the 1e-4 bar; gradients
    are held to 1e-3 or to `mult` x the fp32 reference's own round-off on that tensor (fp32 oracle vs float64 oracle)."""
    o = _oracle("cfg4", need64=True)
    net = _net(o, precision)
    ei = _gpu_edges(o)
    lerr = _layer_errors(o, net, ei)
    for l, (eh, ex, _) in enumerate(lerr):
        assert eh <= 1e-4 and ex <= 1e-4, (l, eh, ex)
    loss, grads = _loss_and_grads(o, net, ei)
    ref = o["grads64"]
    floor = {k: _rel(o["grads32"][k], ref[k]) for k in grads}   # the reference's own fp32 round-off
    gerr = {k: _rel(grads[k], ref[k]) for k in grads}
    worst = max(gerr, key=gerr.get)
    over = {k: (gerr[k], floor[k]) for k in gerr if gerr[k] > 1e-3}
    print(f"full-size cfg4 default init {precision}: loss rel {abs(loss - o['loss64']) / abs(o['loss64']):.2e} (fp32 oracle "
          f"{abs(o['loss32'] - o['loss64']) / abs(o['loss64']):.2e}) worst grad {worst} {gerr[worst]:.2e} (oracle fp32 round-off "
          f"{floor[worst]:.2e}); {len(over)} tensors above 1e-3")
    assert abs(loss - o["loss64"]) <= max(1e-5, mult * abs(o["loss32"] - o["loss64"])) * abs(o["loss64"])
    bad = {k: v for k, v in over.items() if v[0] > mult * v[1]}
    assert not bad, bad


def _fwd_bwd(net, ei, h, x, node_ptr):
    for p in net.parameters():
        p.grad = None
    hc, xc = h.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ho, xo = net(ei, hc, xc, node_ptr=node_ptr)
    (ho.square().sum() + (xo - x).square().sum()).backward()
    grads = [p.grad.clone() for p in net.parameters()] + [hc.grad.clone(), xc.grad.clone()]
    return ho.detach(), xo.detach(), grads


@pytest.mark.parametrize("name", ["cfg2", "cfg4"])
def test_fullsize_rerun_bit_identical(lib, name):
    o = _oracle(name + ("_tamed" if name == "cfg4" else ""))
    net = _net(o, "bf16x3")
    ei = _gpu_edges(o)
    np_ = o["node_ptr"].to(torch.int32).to(DEV)
    h, x = o["h_in"].to(DEV), o["pos_t"].to(DEV)
    h1, x1, g1 = _fwd_bwd(net, ei, h, x, np_)
    h2, x2, g2 = _fwd_bwd(net, ei, h, x, np_)
    assert torch.equal(h1, h2) and torch.equal(x1, x2)
    assert all(torch.equal(a, b) for a, b in zip(g1, g2))
    assert all(bool(torch.isfinite(v).all()) for v in g1)


@pytest.mark.parametrize("name", ["cfg2", "cfg4"])
def test_fullsize_equivariance(lib, name):
    o = _oracle(name + ("_tamed" if name == "cfg4" else ""))
    net = _net(o, "bf16x3")
    ei = _gpu_edges(o)
    np_ = o["node_ptr"].to(torch.int32).to(DEV)
    h, x = o["h_in"].to(DEV), o["pos_t"].to(DEV)
    rng = np.random.default_rng(9)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    R = torch.from_numpy(q).float().to(DEV)
    shift = torch.tensor([0.7, -1.3, 2.1], device=DEV)
    with torch.no_grad():
        h0, x0 = net(ei, h, x, node_ptr=np_)
        x_rot = x @ R.t() + shift
        h1, x1 = net(ei, h, x_rot, node_ptr=np_)
    eh = _rel(h1, h0)
    ex = float(((x1 - x_rot) - (x0 - x) @ R.t()).abs().max() / (x0 - x).abs().max().clamp_min(1e-30))
    print(f"full-size {name}: equivariance h {eh:.2e} dx {ex:.2e}")
    # the rotated coordinates are re-rounded to fp32 (cell side up to 28.5 A) before ||x_i - x_j||^2 is formed
    assert eh <= 2e-3 and ex <= 2e-3, (eh, ex)
