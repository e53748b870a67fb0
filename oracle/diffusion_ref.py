"""TEST INFRASTRUCTURE ONLY (CPU oracle) -- never imported by the product package.

Pure-torch CPU restatement of the reference diffusion maths.  Follows (reference file:line):
  * diffusion_x_h.py:5-14   remove_mean (global mean, or per-graph mean written IN PLACE when batch_index given);
                            identical copy at E3diffusion_new.py:5-14
  * diffusion_x_h.py:92-106 polynomial_schedule + clip_noise_schedule (the array named alphas2 is used AS alpha)
  * diffusion_x_h.py:17-30  alpha_schedule / sigma_schedule = sqrt(1-alpha^2), t = linspace(0,T,T+1)
  * diffusion_x_h.py:51-59  diffuse_zero_to_t (noise COM-free only for mode 'pos')
  * diffusion_x_h.py:61-73  calculate_mu  : mu = z/a_ts - s2_ts*eps/a_ts/sigma_t
  * diffusion_x_h.py:75-90  reverse_diffuse_one_step : z = mu + sqrt(s2_ts*s2_s/s2_t)*noise
  * E3diffusion_new.py:63-98 x-only variant written through x_hat
Noise is always an explicit argument here (the reference draws it with normal_() inside); that is what lets
trajectory-level parity tests inject identical noise on both sides.

Parity pin: outputs of the unmodified reference files imported in the build container
(oracle/make_golden.py), frozen in tests/golden/diffusion_kat.npz.
"""
from __future__ import annotations

import torch


def clip_noise_schedule(alphas2, clip_value=0.001):
    alphas2 = torch.cat([torch.ones(1), alphas2], dim=0)
    step = torch.clamp(alphas2[1:] / alphas2[:-1], min=clip_value, max=1.0)
    return torch.cumprod(step, dim=0)


def polynomial_schedule(T, s=1e-4, power=3.0):
    x = torch.linspace(0, T, T + 1)
    a = torch.pow(1 - torch.pow(x / T, power), 2)
    a = clip_noise_schedule(a, 0.001)
    return (1 - 2 * s) * a + s


class Schedule:
    """alpha/sigma tables of E3DiffusionProcess(s, power, T, 'predefined')."""

    def __init__(self, s, power, T):
        self.T = T
        self.alpha = polynomial_schedule(T, s=s, power=power)
        self.sigma = torch.sqrt(1 - self.alpha ** 2)

    def step_scalars(self, t):
        """(alpha_ts, sigma2_ts, sigma_t, std) exactly as diffusion_x_h.py:61-84 computes them (fp32 tensors)."""
        a_t, a_s = self.alpha[t], self.alpha[t - 1]
        s2_t, s2_s = 1 - a_t ** 2, 1 - a_s ** 2
        a_ts = a_t / a_s
        s2_ts = s2_t - torch.pow(a_ts, 2) * s2_s
        return a_ts, s2_ts, torch.sqrt(s2_t), torch.sqrt(s2_ts * s2_s / s2_t)


def remove_mean(x, batch_index=None):
    if batch_index is None:
        return x - torch.mean(x, dim=0, keepdim=True)
    G = int(batch_index.max()) + 1
    for g in range(G):
        sel = batch_index == g
        x[sel] = x[sel] - torch.mean(x[sel], dim=0, keepdim=True)
    return x


def diffuse_zero_to_t(sch, z, t, noise, mode="pos"):
    """Returns (z_t, noise_used).  `noise` is the raw N(0,1) draw."""
    if mode == "pos":
        noise = remove_mean(noise)
    return sch.alpha[t] * z + sch.sigma[t] * noise, noise


def calculate_mu(sch, z, eps, t):
    a_ts, s2_ts, sig_t, _ = sch.step_scalars(t)
    return z / a_ts - s2_ts * eps / a_ts / sig_t


def reverse_diffuse_one_step(sch, z, eps, t, noise, mode="pos"):
    mu = calculate_mu(sch, z, eps, t)
    std = sch.step_scalars(t)[3]
    if mode == "pos":
        noise = remove_mean(noise)
    return mu + std * noise


# ---- x-only variant (E3diffusion_new.py) -------------------------------------------------------------
def calculate_mu_xonly(sch, pos, eps, t):
    a_t, a_s = sch.alpha[t], sch.alpha[t - 1]
    s2_t, s2_s = 1 - a_t ** 2, 1 - a_s ** 2
    a_ts = a_t / a_s
    s2_ts = s2_t - torch.pow(a_ts, 2) * s2_s
    x_hat = (pos - sch.sigma[t] * eps) / a_t
    return a_ts * s2_s * pos / s2_t + a_s * s2_ts * x_hat / s2_t


def reverse_diffuse_one_step_xonly(sch, mu, t, noise):
    std = sch.step_scalars(t)[3]
    return mu + std * remove_mean(noise)
