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
                 linear_end: float = 0.0205, two_cond: bool = False, first_stage_decoder: nn.Module = None,
                 first_stage_encoder: nn.Module = None, scale_factor: float = 1.0, cond_stage_model_2: nn.Module = None):
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
        # optional first stage (``VQDecoder`` / ``VQEncoder``): latent <-> image like LatentDiffusion (ddpm.py:541-548,
        # 707-771, 827-866; no shipped config sets split_input_params, so the patch-wise branches are not built)
        self.first_stage_decoder = first_stage_decoder
        self.first_stage_encoder = first_stage_encoder
        self.scale_factor = float(scale_factor)
        # talking-face audio conditioning stage (ddpm2cond.py cond_stage_model_2; ``Conv1DTemporalAttention``)
        self.cond_stage_model_2 = cond_stage_model_2

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

    @torch.no_grad()
    def decode_first_stage(self, z, predict_cids=False, force_not_quantize=False):
        """ddpm.py:707-771 (tensor latents): ``first_stage_model.decode(z / scale_factor)``."""
        if self.first_stage_decoder is None:
            raise RuntimeError("DiffusionHandle was built without a first_stage_decoder")
        if predict_cids:
            raise NotImplementedError("predict_cids (logits over code ids) is not used by any shipped config")
        return self.first_stage_decoder.decode(1.0 / self.scale_factor * z, force_not_quantize=force_not_quantize)

    @torch.no_grad()
    def encode_first_stage(self, x):
        """ddpm.py:827-866: ``first_stage_model.encode(x)`` (VQModelInterface: no quantisation)."""
        if self.first_stage_encoder is None:
            raise RuntimeError("DiffusionHandle was built without a first_stage_encoder")
        return self.first_stage_encoder.encode(x)

    def get_first_stage_encoding(self, encoder_posterior):
        """ddpm.py:541-548 for tensor posteriors (VQ first stage)."""
        if not isinstance(encoder_posterior, torch.Tensor):
            raise NotImplementedError(f"encoder_posterior of type '{type(encoder_posterior)}' not yet implemented")
        return self.scale_factor * encoder_posterior
