"""Drop-in for the reference's E3diffusion_new.py (x-only diffusion used by train.py / test.py).

Same schedule as diffusion_x_h; the posterior mean is written through x_hat (E3diffusion_new.py:63-83) and
``reverse_diffuse_one_step(mu, t)`` only adds the COM-free noise (:85-98).
"""
from __future__ import annotations

import torch

from .diffusion_x_h import E3DiffusionProcess as _JointProcess
from .diffusion_x_h import remove_mean


class E3DiffusionProcess(_JointProcess):
    def diffuse_zero_to_t(self, pos, t: int):
        return super().diffuse_zero_to_t(pos, t, mode="pos")

    def calculate_mu(self, pos, epsilon, t: int):
        dev = pos.device
        a_t, a_s = self.alpha(t).to(dev), self.alpha(t - 1).to(dev)
        s2_t, s2_s = 1 - a_t ** 2, 1 - a_s ** 2
        a_ts = a_t / a_s
        s2_ts = s2_t - torch.pow(a_ts, 2) * s2_s
        x_hat = (pos - self.sigma(t).to(dev) * epsilon) / a_t
        return a_ts * s2_s * pos / s2_t + a_s * s2_ts * x_hat / s2_t

    def reverse_diffuse_one_step(self, mu, t: int):
        std = self.step_scalars(t)[3].to(mu.device)
        noise = torch.zeros_like(mu).normal_(mean=0, std=1)
        return mu + std * remove_mean(noise)


__all__ = ["E3DiffusionProcess", "remove_mean"]
