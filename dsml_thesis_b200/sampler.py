"""``DDIMSampler`` — drop-in for the reference samplers

  * ``ldm.models.diffusion.ddim.DDIMSampler``       (face_reenactment/ldm/models/diffusion/ddim.py:11-219)
  * ``ldm.models.diffusion.ddim2cond.DDIMSampler``  (talking_face/.../ddim2cond.py, 2-key dict conditioning;
    face_reenactment/.../ddim2cond.py with its own key names)

Same constructor, ``make_schedule`` / ``sample`` / ``ddim_sampling`` / ``p_sample_ddim`` /
``stochastic_encode`` (/ ``decode``) signatures, same return tuples.  What changes is the per-step update:
the reference's 4x ``torch.full`` + ~11 elementwise launches + 3 hidden device->host reads of the schedule
scalars become ONE fused kernel, ``ldmb200_ddim_step``, fed with host-side fp32 coefficients.
The UNet is reached exactly like in the reference, through ``model.apply_model`` (seam 1).

Random numbers: like the reference, one ``torch.randn(shape, device)`` per step through torch's generator
(util.py:264-267) — also at eta = 0 — so seeded runs consume the RNG stream identically.
"""
from __future__ import annotations

import contextlib
from typing import Optional

import numpy as np
import torch

from . import _cabi
from .unet import UNetModel


# ----------------------------------------------------------------------------------------------
# schedule helpers (host side) — util.py:46-74
# ----------------------------------------------------------------------------------------------

def make_ddim_timesteps(ddim_discr_method, num_ddim_timesteps, num_ddpm_timesteps, verbose=True):
    """util.py:46-60: the DDIM timestep subset, shifted by +1."""
    if ddim_discr_method == "uniform":
        c = num_ddpm_timesteps // num_ddim_timesteps
        ddim_timesteps = np.asarray(list(range(0, num_ddpm_timesteps, c)))
    elif ddim_discr_method == "quad":
        ddim_timesteps = ((np.linspace(0, np.sqrt(num_ddpm_timesteps * .8), num_ddim_timesteps)) ** 2).astype(int)
    else:
        raise NotImplementedError(f'There is no ddim discretization method called "{ddim_discr_method}"')
    steps_out = ddim_timesteps + 1
    if verbose:
        print(f"Selected timesteps for ddim sampler: {steps_out}")
    return steps_out


def make_ddim_sampling_parameters(alphacums, ddim_timesteps, eta, verbose=True):
    """util.py:63-74.  ``alphacums`` is the CPU fp32 tensor of cumulative alphas; the mixed tensor / numpy
    dtypes of the reference expression are kept as they are, they decide the fp32 rounding of sigma."""
    alphas = alphacums[ddim_timesteps]
    alphas_prev = np.asarray([alphacums[0]] + alphacums[ddim_timesteps[:-1]].tolist())
    sigmas = eta * np.sqrt((1 - alphas_prev) / (1 - alphas) * (1 - alphas / alphas_prev))
    if verbose:
        print(f"Selected alphas for ddim sampler: a_t: {alphas}; a_(t-1): {alphas_prev}")
        print(f"For the chosen value of eta, which is {eta}, "
              f"this results in the following sigma_t schedule for ddim sampler {sigmas}")
    return sigmas, alphas, alphas_prev


def _f32(v) -> float:
    """A python float holding the fp32 rounding of ``v`` — what ``torch.full((b,1,1,1), v)`` stores (ddim.py:188-191)."""
    if isinstance(v, torch.Tensor):
        v = v.item()
    return float(np.float32(v))


def _find_unets(model):
    if isinstance(model, torch.nn.Module):
        return [m for m in model.modules() if isinstance(m, UNetModel)]
    return []


class DDIMSampler(object):
    """ddim.py:11-219 (FR flavour: tensor conditioning, CFG by batch doubling)."""

    def __init__(self, model, schedule="linear", **kwargs):
        super().__init__()
        self.model = model
        self.ddpm_num_timesteps = model.num_timesteps
        self.schedule = schedule
        self._coef = None

    # ---- conditioning flavour hooks --------------------------------------------------------------
    def _check_conditioning(self, conditioning, batch_size):
        if conditioning is not None:
            if isinstance(conditioning, dict):
                cbs = conditioning[list(conditioning.keys())[0]].shape[0]
                if cbs != batch_size:
                    print(f"Warning: Got {cbs} conditionings but batch-size is {batch_size}")
            elif conditioning.shape[0] != batch_size:
                print(f"Warning: Got {conditioning.shape[0]} conditionings but batch-size is {batch_size}")

    def _eps(self, x, c, t, scale, uc):
        """ddim.py:170-177 -> (e_cond, e_uncond or None); the CFG mix happens inside the fused update."""
        if uc is None or scale == 1.:
            return self.model.apply_model(x, t, c), None
        x_in = torch.cat([x] * 2)
        t_in = torch.cat([t] * 2)
        c_in = torch.cat([uc, c])
        e_u, e_c = self.model.apply_model(x_in, t_in, c_in).chunk(2)
        return e_c, e_u

    def _finish(self, img, intermediates):
        return img, intermediates

    # ---- schedule ----------------------------------------------------------------------------------
    def register_buffer(self, name, attr):
        # ddim.py:18-22 moves every tensor to "cuda"; here: to the model's device (identical on a GPU box)
        if type(attr) == torch.Tensor:
            dev = self.model.device if hasattr(self.model, "device") else self.model.betas.device
            if attr.device != dev:
                attr = attr.to(dev)
        setattr(self, name, attr)

    def make_schedule(self, ddim_num_steps, ddim_discretize="uniform", ddim_eta=0., verbose=True):
        """ddim.py:24-53."""
        self.ddim_timesteps = make_ddim_timesteps(ddim_discr_method=ddim_discretize,
                                                  num_ddim_timesteps=ddim_num_steps,
                                                  num_ddpm_timesteps=self.ddpm_num_timesteps, verbose=verbose)
        self._set_schedule(ddim_eta, verbose)

    def _set_schedule(self, ddim_eta, verbose):
        """Everything of make_schedule after the timestep subset is chosen (ddim.py:27-53)."""
        alphas_cumprod = self.model.alphas_cumprod
        assert alphas_cumprod.shape[0] == self.ddpm_num_timesteps, 'alphas have to be defined for each timestep'
        to_torch = lambda x: x.clone().detach().to(torch.float32).to(self.model.betas.device)
        ac = alphas_cumprod.detach().cpu()
        self.register_buffer('betas', to_torch(self.model.betas))
        self.register_buffer('alphas_cumprod', to_torch(alphas_cumprod))
        self.register_buffer('alphas_cumprod_prev', to_torch(self.model.alphas_cumprod_prev))
        self.register_buffer('sqrt_alphas_cumprod', to_torch(np.sqrt(ac)))
        self.register_buffer('sqrt_one_minus_alphas_cumprod', to_torch(np.sqrt(1. - ac)))
        self.register_buffer('log_one_minus_alphas_cumprod', to_torch(np.log(1. - ac)))
        self.register_buffer('sqrt_recip_alphas_cumprod', to_torch(np.sqrt(1. / ac)))
        self.register_buffer('sqrt_recipm1_alphas_cumprod', to_torch(np.sqrt(1. / ac - 1)))
        ddim_sigmas, ddim_alphas, ddim_alphas_prev = make_ddim_sampling_parameters(
            alphacums=ac, ddim_timesteps=self.ddim_timesteps, eta=ddim_eta, verbose=verbose)
        sqrt_1ma = np.sqrt(1. - ddim_alphas)
        # host-side per-step scalars, rounded to fp32 like torch.full does: no device->host read in the loop
        n = len(self.ddim_timesteps)
        sqrt_1ma_prev = np.sqrt(1. - ddim_alphas_prev)   # ddim2.py:96 (float64 numpy, like ddim_alphas_prev)
        self._coef = dict(a_t=[_f32(ddim_alphas[i]) for i in range(n)],
                          a_prev=[_f32(ddim_alphas_prev[i]) for i in range(n)],
                          sigma=[_f32(ddim_sigmas[i]) for i in range(n)],
                          s1=[_f32(sqrt_1ma[i]) for i in range(n)],
                          s1_prev=[_f32(sqrt_1ma_prev[i]) for i in range(n)])
        self.register_buffer('ddim_sqrt_one_minus_alphas_prev', sqrt_1ma_prev)
        self._coef_orig = None
        self.register_buffer('ddim_sigmas', ddim_sigmas)
        self.register_buffer('ddim_alphas', ddim_alphas)
        self.register_buffer('ddim_alphas_prev', ddim_alphas_prev)
        self.register_buffer('ddim_sqrt_one_minus_alphas', sqrt_1ma)
        acp = self.model.alphas_cumprod_prev.detach().cpu().to(torch.float32)
        acf = ac.to(torch.float32)
        sig_orig = ddim_eta * torch.sqrt((1 - acp) / (1 - acf) * (1 - acf / acp))
        self.register_buffer('ddim_sigmas_for_original_num_steps', sig_orig)

    def _original_coef(self):
        if self._coef_orig is None:
            ac = self.model.alphas_cumprod.detach().cpu().float()
            acp = self.model.alphas_cumprod_prev.detach().cpu().float()
            s1 = self.model.sqrt_one_minus_alphas_cumprod.detach().cpu().float() \
                if hasattr(self.model, "sqrt_one_minus_alphas_cumprod") else torch.sqrt(1. - ac)
            sg = self.ddim_sigmas_for_original_num_steps.detach().cpu().float()
            self._coef_orig = dict(a_t=ac.tolist(), a_prev=acp.tolist(), sigma=sg.tolist(), s1=s1.tolist())
        return self._coef_orig

    # ---- sampling ----------------------------------------------------------------------------------
    @torch.no_grad()
    def sample(self, S, batch_size, shape, conditioning=None, callback=None, normals_sequence=None,
               img_callback=None, quantize_x0=False, eta=0., mask=None, x0=None, temperature=1.,
               noise_dropout=0., score_corrector=None, corrector_kwargs=None, verbose=True, x_T=None,
               log_every_t=100, unconditional_guidance_scale=1., unconditional_conditioning=None, **kwargs):
        """ddim.py:56-110."""
        self._check_conditioning(conditioning, batch_size)
        self.make_schedule(ddim_num_steps=S, ddim_eta=eta, verbose=verbose)
        C, H, W = shape
        size = (batch_size, C, H, W)
        if verbose:
            print(f'Data shape for DDIM sampling is {size}, eta {eta}')
        return self.ddim_sampling(conditioning, size, callback=callback, img_callback=img_callback,
                                  quantize_denoised=quantize_x0, mask=mask, x0=x0,
                                  ddim_use_original_steps=False, noise_dropout=noise_dropout,
                                  temperature=temperature, score_corrector=score_corrector,
                                  corrector_kwargs=corrector_kwargs, x_T=x_T, log_every_t=log_every_t,
                                  unconditional_guidance_scale=unconditional_guidance_scale,
                                  unconditional_conditioning=unconditional_conditioning, verbose=verbose)

    @torch.no_grad()
    def ddim_sampling(self, cond, shape, x_T=None, ddim_use_original_steps=False, callback=None, timesteps=None,
                      quantize_denoised=False, mask=None, x0=None, img_callback=None, log_every_t=100,
                      temperature=1., noise_dropout=0., score_corrector=None, corrector_kwargs=None,
                      unconditional_guidance_scale=1., unconditional_conditioning=None, verbose=False):
        """ddim.py:113-162."""
        device = self.model.betas.device
        b = shape[0]
        img = torch.randn(shape, device=device) if x_T is None else x_T
        if timesteps is None:
            timesteps = self.ddpm_num_timesteps if ddim_use_original_steps else self.ddim_timesteps
        elif not ddim_use_original_steps:
            subset_end = int(min(timesteps / self.ddim_timesteps.shape[0], 1) * self.ddim_timesteps.shape[0]) - 1
            timesteps = self.ddim_timesteps[:subset_end]
        intermediates = {'x_inter': [img], 'pred_x0': [img]}
        time_range = reversed(range(0, timesteps)) if ddim_use_original_steps else np.flip(timesteps)
        total_steps = timesteps if ddim_use_original_steps else timesteps.shape[0]
        if verbose:
            print(f"Running DDIM Sampling with {total_steps} timesteps")
        with contextlib.ExitStack() as stack:
            for u in _find_unets(self.model):   # weights cannot change inside the loop: pack them once
                stack.enter_context(u.frozen_weights())
            for i, step in enumerate(time_range):
                index = total_steps - i - 1
                ts = torch.full((b,), int(step), device=device, dtype=torch.long)
                if mask is not None:
                    assert x0 is not None
                    img_orig = self.model.q_sample(x0, ts)
                    img = img_orig * mask + (1. - mask) * img
                img, pred_x0 = self.p_sample_ddim(
                    img, cond, ts, index=index, use_original_steps=ddim_use_original_steps,
                    quantize_denoised=quantize_denoised, temperature=temperature, noise_dropout=noise_dropout,
                    score_corrector=score_corrector, corrector_kwargs=corrector_kwargs,
                    unconditional_guidance_scale=unconditional_guidance_scale,
                    unconditional_conditioning=unconditional_conditioning)
                if callback:
                    callback(i)
                if img_callback:
                    img_callback(pred_x0, i)
                if index % log_every_t == 0 or index == total_steps - 1:
                    intermediates['x_inter'].append(img)
                    intermediates['pred_x0'].append(pred_x0)
        return self._finish(img, intermediates)

    @torch.no_grad()
    def p_sample_ddim(self, x, c, t, index, repeat_noise=False, use_original_steps=False, quantize_denoised=False,
                      temperature=1., noise_dropout=0., score_corrector=None, corrector_kwargs=None,
                      unconditional_guidance_scale=1., unconditional_conditioning=None):
        """ddim.py:165-203: one (CFG: batch-doubled) UNet eval, then the fused x_{t-1} update (K13)."""
        if not x.is_cuda:
            raise _cabi.Ldmb200Error("dsml_thesis_b200.DDIMSampler runs on a CUDA device only (no CPU fallback)")
        e_c, e_u = self._eps(x, c, t, unconditional_guidance_scale, unconditional_conditioning)
        if score_corrector is not None:
            assert self.model.parameterization == "eps"
            e_mix = e_c if e_u is None else e_u + unconditional_guidance_scale * (e_c - e_u)
            e_c, e_u = score_corrector.modify_score(self.model, e_mix, x, t, c, **(corrector_kwargs or {})), None
        k = self._original_coef() if use_original_steps else self._coef
        a_t, a_prev, sigma, s1 = k["a_t"][index], k["a_prev"][index], k["sigma"][index], k["s1"][index]
        shape = x.shape
        if repeat_noise:  # util.py:264-267
            noise = torch.randn((1, *shape[1:]), device=x.device).repeat(shape[0], *((1,) * (len(shape) - 1)))
        else:
            noise = torch.randn(shape, device=x.device)
        late_noise = None
        if noise_dropout > 0. or quantize_denoised:
            late_noise = float(np.float32(sigma)) * noise * temperature
            if noise_dropout > 0.:
                late_noise = torch.nn.functional.dropout(late_noise, p=noise_dropout)
            noise = None
        x_prev, pred_x0 = ddim_step(x, e_c, e_u, noise, a_t, a_prev, sigma, s1,
                                    float(unconditional_guidance_scale), float(temperature))
        if quantize_denoised:
            # rarely used branch (ddim.py:195-196): x_prev must be rebuilt from the QUANTISED pred_x0
            e_t = e_c if e_u is None else e_u + unconditional_guidance_scale * (e_c - e_u)
            pred_x0, _, *_ = self.model.first_stage_model.quantize(pred_x0)
            f = lambda v: torch.full((shape[0], 1, 1, 1), v, device=x.device)
            x_prev = f(a_prev).sqrt() * pred_x0 + (1. - f(a_prev) - f(sigma) ** 2).sqrt() * e_t + late_noise
        elif late_noise is not None:
            x_prev = x_prev + late_noise
        return x_prev, pred_x0

    @torch.no_grad()
    def stochastic_encode(self, x0, t, use_original_steps=False, noise=None):
        """ddim.py:206-219 (with ``extract_into_tensor`` = gather + reshape, util.py:96-99)."""
        if use_original_steps:
            sqrt_alphas_cumprod = self.sqrt_alphas_cumprod
            sqrt_one_minus_alphas_cumprod = self.sqrt_one_minus_alphas_cumprod
        else:
            sqrt_alphas_cumprod = torch.sqrt(torch.as_tensor(self.ddim_alphas))
            sqrt_one_minus_alphas_cumprod = torch.as_tensor(self.ddim_sqrt_one_minus_alphas)
        if noise is None:
            noise = torch.randn_like(x0)
        ex = lambda a: a.to(x0.device).gather(-1, t).reshape(t.shape[0], *((1,) * (x0.dim() - 1)))
        return ex(sqrt_alphas_cumprod) * x0 + ex(sqrt_one_minus_alphas_cumprod) * noise


class DDIMSampler2Cond(DDIMSampler):
    """talking_face/ldm/models/diffusion/ddim2cond.py: ``conditioning`` is a dict with exactly two keys, passed
    to ``model.apply_model(x, t, c12, c34)`` as two positionals (ddim2cond.py:80-81,165-169); CFG runs two
    evals at batch B; ``sample`` returns ``(samples, x_inter list)`` (ddim2cond.py:156)."""

    CROSS_KEY = 'class_label_&_audio'
    CONCAT_KEY = 'motion_&_id'
    RETURN_DICT = False

    def _check_conditioning(self, conditioning, batch_size):
        assert conditioning is not None
        assert len(conditioning.keys()) == 2

    def _eps(self, x, c, t, scale, uc):
        e_c = self.model.apply_model(x, t, c[self.CROSS_KEY], c[self.CONCAT_KEY])
        if uc is None or scale == 1.:
            return e_c, None
        return e_c, self.model.apply_model(x, t, uc, c[self.CONCAT_KEY])

    def _finish(self, img, intermediates):
        return (img, intermediates) if self.RETURN_DICT else (img, intermediates['x_inter'])

    def ddim_sampling(self, *a, log_every_t=50, **k):
        return super().ddim_sampling(*a, log_every_t=log_every_t, **k)

    @torch.no_grad()
    def decode(self, x_latent, cond, t_start, unconditional_guidance_scale=1.0, unconditional_conditioning=None,
               use_original_steps=False):
        """ddim2cond.py:230-250."""
        timesteps = np.arange(self.ddpm_num_timesteps) if use_original_steps else self.ddim_timesteps
        timesteps = timesteps[:t_start]
        total_steps = timesteps.shape[0]
        x_dec = x_latent
        for i, step in enumerate(np.flip(timesteps)):
            index = total_steps - i - 1
            ts = torch.full((x_latent.shape[0],), int(step), device=x_latent.device, dtype=torch.long)
            x_dec, _ = self.p_sample_ddim(x_dec, cond, ts, index=index, use_original_steps=use_original_steps,
                                          unconditional_guidance_scale=unconditional_guidance_scale,
                                          unconditional_conditioning=unconditional_conditioning)
        return x_dec


    @torch.no_grad()
    def progressive_sampling(self, c1, xid, xmasks, audio_feats, S, batch_size, num_frames, shape, audio_window, eta=0.,
                             verbose=True, unconditional_guidance_scale=1., unconditional_conditioning=None, x_T=None,
                             **kwargs):
        """The autoregressive talking-face sampler of TF/progressive_sampling_difftalk.py:246-319 (the method of that
        script's private sampler class, same arguments and return value): per frame, the audio window goes through
        ``model.cond_stage_model_2``, the masked frame through ``model.encode_first_stage``, the frame is generated by
        an S-step DDIM loop from fresh noise, and its latent becomes the reference (identity) latent of the next frame
        (``zid = img``, :316-317).  Returns ``(generated_frame_latents, None)``.

        Extensions (not in the reference, which is batch 1): ``x_T`` — a tensor ``[frames, B, C, H, W]`` (or a list) of
        starting noise instead of the per-frame ``torch.randn`` draw (parity tests); and B independent streams at once:
        ``c1 [B,1,256]``, ``xid [B,c,h,w]``, ``xmasks [frames,B,3,H,W]``, ``audio_feats [B,frames,768]``.
        The reference's CFG branch concatenates on ``dim=21`` (:299) and cannot run; here it is ``dim=2``."""
        assert c1 is not None
        assert eta in [0., 1.]
        self.make_schedule(ddim_num_steps=S, ddim_eta=eta, verbose=verbose)
        if audio_feats.dim() == 3 and audio_feats.shape[0] == 1:
            audio_feats = audio_feats.squeeze(0)
        streams = audio_feats.dim() == 3                      # [B, frames, 768]
        num_frames = audio_feats.shape[1] if streams else audio_feats.shape[0]
        C, H, W = shape
        size = (batch_size, C, H, W)
        if verbose:
            print(f'Data shape for DDIM sampling is {size}, eta {eta}')
        device = self.model.betas.device
        zid = xid.clone().detach()
        generated_frame_latents = list()
        for frame_idx in range(num_frames):
            audio_indices = [min(max(frame_idx + i, 0), num_frames - 1) for i in range(-audio_window, audio_window + 1)]
            assert audio_indices[audio_window] == frame_idx
            c2 = audio_feats[:, audio_indices] if streams else audio_feats[audio_indices].unsqueeze(0)
            c2 = self.model.cond_stage_model_2(c2)                               # (B, 1, 768)
            c12 = torch.cat([c1, c2], dim=2)                                     # (B, 1, 1024)
            masked_img = xmasks[frame_idx]
            if masked_img.dim() == 3:
                masked_img = masked_img.unsqueeze(0)
            c3 = self.model.encode_first_stage(masked_img)                       # masked frame latent (B, c, h, w)
            c34 = torch.cat([c3, zid], dim=1)                                    # (B, 2c, h, w)
            uc = torch.cat([unconditional_conditioning, c2], dim=2) if unconditional_conditioning is not None else None
            cond = {self.CROSS_KEY: c12, self.CONCAT_KEY: c34}
            img = torch.randn(size, device=device) if x_T is None else x_T[frame_idx].to(device)
            img, _ = self.ddim_sampling(cond, size, x_T=img, unconditional_guidance_scale=unconditional_guidance_scale,
                                        unconditional_conditioning=uc, verbose=False)
            generated_frame_latents.append(img)
            zid = img.clone().detach()
        return generated_frame_latents, None


class DDIMSampler2CondFR(DDIMSampler2Cond):
    """face_reenactment/ldm/models/diffusion/ddim2cond.py: keys ``class_label`` / ``shape_image`` (line 166),
    ``sample`` returns the intermediates dict (line 157)."""
    CROSS_KEY = 'class_label'
    CONCAT_KEY = 'shape_image'
    RETURN_DICT = True

    def ddim_sampling(self, *a, log_every_t=100, **k):
        return DDIMSampler.ddim_sampling(self, *a, log_every_t=log_every_t, **k)


def make_ddim_timesteps_strength(ddim_discr_method, num_ddim_timesteps, num_ddpm_timesteps, strength=1.0, verbose=True):
    """ddim2.py:9-29 / latent_manipulation.py:52-72: S timesteps spread over [0, T*strength], first one forced to 1
    (no +1 shift) — the grid of the DDIM inversion / re-generation used for emotion manipulation."""
    if ddim_discr_method == 'uniform':
        ddim_timesteps = np.linspace(0, 1, num_ddim_timesteps) * int(num_ddpm_timesteps * strength)
        ddim_timesteps = [int(s) for s in list(ddim_timesteps)]
    elif ddim_discr_method == 'quad':
        ddim_timesteps = ((np.linspace(0, np.sqrt(num_ddpm_timesteps * .8), num_ddim_timesteps)) ** 2).astype(int)
        ddim_timesteps = [int(s) for s in list(ddim_timesteps)]
    else:
        raise NotImplementedError(f'There is no ddim discretization method called "{ddim_discr_method}"')
    steps_out = np.asarray([1] + ddim_timesteps[1:])
    if verbose:
        print(f'Selected timesteps for ddim sampler: {steps_out}')
    return steps_out


class DDIMSampler2(DDIMSampler):
    """``ldm.models.diffusion.ddim2.DDIMSampler2`` (FR ddim2.py:53-284) plus the DDIM inversion that the thesis'
    manipulation scripts carry as private sampler copies (``q_sample_ddim`` / ``latent_manipulation``,
    FR/latent_manipulation.py:377-489; ``compute_latents.py:257``): SURVEY §8f row n1.  The inversion step is the
    SAME fused update kernel with the roles of (a_t, a_prev) swapped and sigma = 0, so it is bit-identical to the
    reference's chain of torch ops like the sampling step."""

    def make_schedule(self, ddim_num_steps, ddim_discretize="uniform", ddim_eta=0., strength=1.0, verbose=True):
        """ddim2.py:66-97."""
        self.ddim_timesteps = make_ddim_timesteps_strength(ddim_discretize, ddim_num_steps, self.ddpm_num_timesteps,
                                                           strength=strength, verbose=verbose)
        self._set_schedule(ddim_eta, verbose)

    @torch.no_grad()
    def q_sample_ddim(self, x, c, t, index, repeat_noise=False, use_original_steps=False, quantize_denoised=False,
                      temperature=1., noise_dropout=0., score_corrector=None, corrector_kwargs=None,
                      unconditional_guidance_scale=1., unconditional_conditioning=None):
        """latent_manipulation.py:377-419: one deterministic step x_t -> x_{t+1} (towards noise)."""
        if not x.is_cuda:
            raise _cabi.Ldmb200Error("dsml_thesis_b200.DDIMSampler2 runs on a CUDA device only (no CPU fallback)")
        if use_original_steps:
            raise NotImplementedError("q_sample_ddim(use_original_steps=True) reads model.sqrt_one_minus_alphas_cumprod_prev, "
                                      "which the reference model does not define either")
        e_c, e_u = self._eps(x, c, t, unconditional_guidance_scale, unconditional_conditioning)
        if score_corrector is not None:
            assert self.model.parameterization == "eps"
            e_mix = e_c if e_u is None else e_u + unconditional_guidance_scale * (e_c - e_u)
            e_c, e_u = score_corrector.modify_score(self.model, e_mix, x, t, c, **(corrector_kwargs or {})), None
        k = self._coef
        at, at_next, s1 = k["a_prev"][index], k["a_t"][index], k["s1_prev"][index]
        if quantize_denoised:
            e_t = e_c if e_u is None else e_u + unconditional_guidance_scale * (e_c - e_u)
            f = lambda v: torch.full((x.shape[0], 1, 1, 1), v, device=x.device)
            pred_x0 = (x - f(s1) * e_t) / f(at).sqrt()
            pred_x0, _, *_ = self.model.first_stage_model.quantize(pred_x0)
            return f(at_next).sqrt() * pred_x0 + (1. - f(at_next)).sqrt() * e_t, pred_x0
        # x_next = sqrt(a_next) * pred_x0 + sqrt(1 - a_next) * e   ==  the sampling update with sigma = 0
        return ddim_step(x, e_c, e_u, None, at, at_next, 0.0, s1, float(unconditional_guidance_scale), 1.0)

    @torch.no_grad()
    def encode(self, x0, c, S, strength=1.0, unconditional_guidance_scale=1., unconditional_conditioning=None,
               verbose=False):
        """DDIM inversion x0 -> x_T' over the strength-scaled grid (compute_latents.py / latent_manipulation.py:470-477)."""
        self.make_schedule(ddim_num_steps=S, ddim_eta=0., verbose=verbose, strength=strength)
        b, device = x0.shape[0], x0.device
        x = x0.clone().detach()
        with contextlib.ExitStack() as stack:
            for u in _find_unets(self.model):
                stack.enter_context(u.frozen_weights())
            for i, step in enumerate(self.ddim_timesteps):
                ts = torch.full((b,), int(step), device=device, dtype=torch.long)
                x, _ = self.q_sample_ddim(x, c, ts, index=i, unconditional_guidance_scale=unconditional_guidance_scale,
                                          unconditional_conditioning=unconditional_conditioning)
        return x

    @torch.no_grad()
    def latent_manipulation(self, c_src, c_trg, S, batch_size, shape, callback=None, normals_sequence=None,
                            img_callback=None, quantize_x0=False, eta=0., mask=None, x0=None, temperature=1.,
                            noise_dropout=0., score_corrector=None, corrector_kwargs=None, verbose=True, x_T=None,
                            log_every_t=100, unconditional_guidance_scale=1., unconditional_conditioning=None,
                            strength=1.0, **kwargs):
        """latent_manipulation.py:421-489: invert x0 under the source condition, regenerate under the target one.
        Returns ``(img, x_T', x0)``."""
        assert c_src is not None and c_trg is not None
        assert x0 is not None
        assert x_T is None
        assert eta == 0
        x = self.encode(x0, c_src, S, strength=strength, unconditional_guidance_scale=unconditional_guidance_scale,
                        unconditional_conditioning=unconditional_conditioning, verbose=verbose)
        b, device = x0.shape[0], x0.device
        timesteps = self.ddim_timesteps
        total_steps = timesteps.shape[0]
        img = x.clone().detach()
        with contextlib.ExitStack() as stack:
            for u in _find_unets(self.model):
                stack.enter_context(u.frozen_weights())
            for i, step in enumerate(np.flip(timesteps)):
                index = total_steps - i - 1
                ts = torch.full((b,), int(step), device=device, dtype=torch.long)
                img, _ = self.p_sample_ddim(img, c_trg, ts, index=index,
                                            unconditional_guidance_scale=unconditional_guidance_scale,
                                            unconditional_conditioning=unconditional_conditioning)
        return img, x, x0


# ----------------------------------------------------------------------------------------------
# K13: the fused update
# ----------------------------------------------------------------------------------------------

def ddim_step(x: torch.Tensor, e_cond: torch.Tensor, e_uncond: Optional[torch.Tensor],
              noise: Optional[torch.Tensor], a_t: float, a_prev: float, sigma_t: float,
              sqrt_one_minus_at: float, cfg_scale: float = 1.0, temperature: float = 1.0):
    """One launch of ``ldmb200_ddim_step`` (ddim.py:177,188-202).  Returns ``(x_prev, pred_x0)``."""
    def prep(t):
        if t is None:
            return None
        if t.dtype != torch.float32 or not t.is_contiguous():
            t = t.float().contiguous()
        if t.data_ptr() % 16:  # e.g. the second half of a chunk(2) view
            t = t.clone()
        return t
    x, e_cond, e_uncond, noise = prep(x), prep(e_cond), prep(e_uncond), prep(noise)
    x_prev, pred_x0 = torch.empty_like(x), torch.empty_like(x)
    with torch.cuda.device(x.device):
        _cabi.call("ldmb200_ddim_step", x.data_ptr(), e_cond.data_ptr(),
                   e_uncond.data_ptr() if e_uncond is not None else None,
                   noise.data_ptr() if noise is not None else None, x_prev.data_ptr(), pred_x0.data_ptr(),
                   x.numel(), a_t, a_prev, sigma_t, sqrt_one_minus_at, cfg_scale, temperature,
                   torch.cuda.current_stream().cuda_stream)
    return x_prev, pred_x0
