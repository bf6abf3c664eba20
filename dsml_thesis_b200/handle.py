"""``DiffusionHandle`` — the minimum of the reference's ``LatentDiffusion`` a sampler needs.

The reference's Lightning modules (ddpm.py, ddpm2cond.py) need pytorch_lightning / omegaconf / a VQGAN
checkpoint; on a box that has them the real ``LatentDiffusion`` is the model handle and this class is not used.
Benchmarks, smoke tests and the multi-GPU sampler use this stand-in: the schedule buffers of
``DDPM.register_schedule`` (ddpm.py:117-142, "linear" = ``linspace(sqrt(s), sqrt(e), T)**2`` in float64,
util.py:22-25) and the ``apply_model`` -> ``DiffusionWrapper.forward`` call boundary
(ddpm.py:893-902,989,1404-1412; talking_face ddpm2cond.py:911-945,1307-1311).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn


class DiffusionHandle(nn.Module):
    def __init__(self, unet: nn.Module, timesteps: int = 1000, linear_start: float = 0.0015,
                 linear_end: float = 0.0205, two_cond: bool = False):
        super().__init__()
        betas = (torch.linspace(linear_start ** 0.5, linear_end ** 0.5, timesteps, dtype=torch.float64) ** 2).numpy()
        alphas_cumprod = np.cumprod(1.0 - betas, axis=0)
        alphas_cumprod_prev = np.append(1.0, alphas_cumprod[:-1])
        f32 = lambda a: torch.tensor(a, dtype=torch.float32)
        self.num_timesteps = int(timesteps)
        self.parameterization = "eps"
        self.register_buffer("betas", f32(betas))
        self.register_buffer("alphas_cumprod", f32(alphas_cumprod))
        self.register_buffer("alphas_cumprod_prev", f32(alphas_cumprod_prev))
        self.register_buffer("sqrt_alphas_cumprod", f32(np.sqrt(alphas_cumprod)))
        self.register_buffer("sqrt_one_minus_alphas_cumprod", f32(np.sqrt(1.0 - alphas_cumprod)))
        self.diffusion_model = unet
        self.two_cond = two_cond

    @property
    def device(self):
        return self.betas.device

    def apply_model(self, x, t, c, c_concat=None):
        """FR: ``unet(x, t, context=c)`` (conditioning_key 'crossattn'); TF: ``unet(cat([x, c34], 1), t, context=c12)``."""
        if self.two_cond:
            return self.diffusion_model(torch.cat([x, c_concat], dim=1), t, context=c)
        return self.diffusion_model(x, t, context=c)

    def q_sample(self, x_start, t, noise=None):
        """ddpm.py:273-276."""
        noise = torch.randn_like(x_start) if noise is None else noise
        e = lambda a: a.gather(-1, t).reshape(-1, 1, 1, 1)
        return e(self.sqrt_alphas_cumprod) * x_start + e(self.sqrt_one_minus_alphas_cumprod) * noise
