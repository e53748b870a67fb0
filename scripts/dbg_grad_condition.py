"""Conditioning of the full-size cfg4 weight gradients: relative change of every gradient tensor (max-abs / max-abs) when
the inputs are perturbed by 1e-6 relative, in the fp32 mode, next to the bf16x3-vs-fp32 difference.
python scripts/dbg_grad_condition.py [cfg4|cfg2]"""
import sys

import torch

sys.path.insert(0, ".")
from tests.test_gpu_fullsize import _fwd_bwd, _rel, _workload  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg4"
net, ei, h, x, node_ptr = _workload(name)
_, _, g0 = _fwd_bwd(net, ei, h, x, node_ptr, "fp32")
gen = torch.Generator().manual_seed(1)
hp = h * (1 + 1e-6 * torch.randn(h.shape, generator=gen).to(h.device))
xp = x * (1 + 1e-6 * torch.randn(x.shape, generator=gen).to(x.device))
_, _, g1 = _fwd_bwd(net, ei, hp, xp, node_ptr, "fp32")
_, _, g3 = _fwd_bwd(net, ei, h, x, node_ptr, "bf16x3")
rows = sorted(((k, _rel(g3[k], g0[k]), _rel(g1[k], g0[k])) for k in g0), key=lambda r: -r[1])
for k, e3, ep in rows[:12]:
    print(f"{k:40s} bf16x3-fp32 {e3:.2e}   fp32(1e-6 perturbed)-fp32 {ep:.2e}   amplification {ep / 1e-6:.0f}x")
