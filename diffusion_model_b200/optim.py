"""Fused Adam on the flat parameter / gradient buffers of a kernel-backed EGNN (one launch per step instead of one per
tensor).  Arithmetic of torch.optim.Adam, which the reference selects with optimizer: 'Adam' (parts/def_for_main.py:133-134).
Any torch optimizer built from ``egnn.parameters()`` keeps working too; this one is the fast path."""
from __future__ import annotations

import torch

from ._lib import call, ptr, stream_ptr


def flat_grad_of(module):
    """The gradient of every parameter of `module` as ONE contiguous fp32 tensor in flat-parameter order.

    The backward kernel writes all parameter gradients into one buffer and hands autograd views of it; when autograd
    kept those views as ``p.grad`` the buffer itself is returned (zero copy), otherwise the grads are concatenated."""
    fp = module._flat
    flat = fp.ensure()
    g = module._last_grad_flat
    if g is not None and g.numel() == flat.numel():
        off, ok = 0, True
        for p in fp.params:
            if p.grad is None or p.grad.data_ptr() != g.data_ptr() + 4 * off:
                ok = False
                break
            off += p.numel()
        if ok:
            return g, True
    parts = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in fp.params]
    return torch.cat(parts), False


def scatter_flat_grad(module, g):
    """Write a flat gradient back into the per-parameter ``.grad`` tensors (used after an all-reduce of a copy)."""
    off = 0
    for p in module._flat.params:
        n = p.numel()
        if p.grad is None:
            p.grad = g[off:off + n].view(p.shape)
        else:
            p.grad.copy_(g[off:off + n].view(p.shape))
        off += n


class FlatAdam:
    def __init__(self, module, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        self.module, self.lr, self.betas, self.eps, self.weight_decay = module, lr, betas, eps, weight_decay
        self.step_count = 0
        self.exp_avg = self.exp_avg_sq = None

    def zero_grad(self, set_to_none=True):
        for p in self.module._flat.params:
            p.grad = None
        self.module._last_grad_flat = None

    def train(self):
        pass

    def eval(self):
        pass

    def step(self, grad_scale=1.0):
        flat = self.module.flat_parameters()
        g, _ = flat_grad_of(self.module)
        if self.exp_avg is None or self.exp_avg.numel() != flat.numel() or self.exp_avg.device != flat.device:
            self.exp_avg = torch.zeros_like(flat)
            self.exp_avg_sq = torch.zeros_like(flat)
        self.step_count += 1
        call("egnn_adam_step", ptr(flat), ptr(g), ptr(self.exp_avg), ptr(self.exp_avg_sq), flat.numel(), float(self.lr),
             float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay), self.step_count,
             float(grad_scale), stream_ptr(flat.device))
        self.module._flat.packed_key = None  # parameters changed behind autograd's back: force a re-pack

    def state_dict(self):
        return dict(step=self.step_count, exp_avg=self.exp_avg, exp_avg_sq=self.exp_avg_sq, lr=self.lr, betas=self.betas,
                    eps=self.eps, weight_decay=self.weight_decay)

    def load_state_dict(self, sd):
        self.step_count, self.exp_avg, self.exp_avg_sq = sd["step"], sd["exp_avg"], sd["exp_avg_sq"]
        self.lr, self.betas, self.eps, self.weight_decay = sd["lr"], sd["betas"], sd["eps"], sd["weight_decay"]
