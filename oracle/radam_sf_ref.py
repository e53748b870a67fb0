"""TEST INFRASTRUCTURE ONLY (CPU oracle) -- never imported by the product package.

``schedulefree.RAdamScheduleFree`` restated per tensor with plain torch ops.  The reference selects it with
``optimizer: RAdamScheduleFree`` (parameters.yaml) through ``RAdamScheduleFree(param_list, lr=lr)``
(parts/def_for_main.py:137-138) and switches it with ``optimizer.train()`` / ``optimizer.eval()`` in the epoch loops
(parts/train_per_iretation.py:112-113, :198-199).

PARITY UNPINNED: the package is a third-party dependency that is neither vendored in the reference checkout nor pinned
by any requirements file there, and it is not installed in this image (no network).  What follows restates its published
algorithm (facebookresearch/schedule_free, ``schedulefree/radam_schedulefree.py``, the version with the
``silent_sgd_phase`` argument): defaults lr=0.0025, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, r=0,
weight_lr_power=2, silent_sgd_phase=True.  There is no golden vector from the package itself; the tests check the CUDA
kernel against THIS restatement and the algorithm's invariants (eval()/train() are inverse maps, the first steps with
rho_t <= 4 are silent, x is the weighted average of the z sequence).
"""
from __future__ import annotations

import torch


class RAdamScheduleFreeRef:
    def __init__(self, params, lr=0.0025, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, r=0.0, weight_lr_power=2.0,
                 silent_sgd_phase=True):
        self.params = list(params)
        self.g = dict(lr=lr, betas=betas, eps=eps, r=r, k=0, train_mode=False, weight_sum=0.0, lr_max=-1.0,
                      scheduled_lr=0.0, weight_lr_power=weight_lr_power, weight_decay=weight_decay,
                      silent_sgd_phase=silent_sgd_phase)
        self.state = {}

    @torch.no_grad()
    def eval(self):
        if self.g["train_mode"]:
            beta1 = self.g["betas"][0]
            for p in self.params:
                if p in self.state:
                    p.lerp_(end=self.state[p]["z"], weight=1 - 1 / beta1)  # p <- x
            self.g["train_mode"] = False

    @torch.no_grad()
    def train(self):
        if not self.g["train_mode"]:
            beta1 = self.g["betas"][0]
            for p in self.params:
                if p in self.state:
                    p.lerp_(end=self.state[p]["z"], weight=1 - beta1)  # p <- y
            self.g["train_mode"] = True

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    @torch.no_grad()
    def step(self):
        g = self.g
        if not g["train_mode"]:
            raise Exception("Optimizer was not in train mode when step is called.")
        beta1, beta2 = g["betas"]
        step = g["k"] + 1
        beta2_t = beta2 ** step
        bias_correction2 = 1 - beta2_t
        rho_inf = 2 / (1 - beta2) - 1
        rho_t = rho_inf - 2 * step * beta2_t / bias_correction2
        rect = (((rho_t - 4) * (rho_t - 2) * rho_inf / ((rho_inf - 4) * (rho_inf - 2) * rho_t)) ** 0.5 if rho_t > 4.0
                else float(not g["silent_sgd_phase"]))
        lr = g["lr"] * rect
        g["scheduled_lr"] = lr
        lr_max = g["lr_max"] = max(lr, g["lr_max"])
        weight = (step ** g["r"]) * (lr_max ** g["weight_lr_power"])
        weight_sum = g["weight_sum"] = g["weight_sum"] + weight
        try:
            ckp1 = weight / weight_sum
        except ZeroDivisionError:
            ckp1 = 0
        adaptive_y_lr = lr * (beta1 * (1 - ckp1) - 1)
        for p in self.params:
            if p.grad is None:
                continue
            if p not in self.state:
                self.state[p] = dict(z=p.detach().clone(), exp_avg_sq=torch.zeros_like(p))
            st = self.state[p]
            grad, z, v = p.grad.clone(), st["z"], st["exp_avg_sq"]
            v.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
            if rho_t > 4.0:
                denom = v.div(bias_correction2).sqrt_().add_(g["eps"])
                grad = grad.div_(denom)
            if g["weight_decay"] != 0:
                grad.add_(p, alpha=g["weight_decay"])  # weight decay evaluated at y
            p.lerp_(end=z, weight=ckp1)
            p.add_(grad, alpha=adaptive_y_lr)
            z.sub_(grad, alpha=lr)
        g["k"] = step
