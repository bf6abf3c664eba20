"""Import shim for the REAL reference (TEST INFRASTRUCTURE ONLY; build container only).

``/root/reference`` exists only in the build container, never on the GPU box, so this module is
used by ``tests/golden/make_golden.py`` (to produce committed golden vectors), by the
``-m "not gpu"`` tests that pin the oracle restatement against the real reference when it is
present, and by ``bench.py --impl reference`` when it is present.  Nothing in the product
package imports it.

What it does:
  * puts ``/root/reference/face_reenactment`` (or ``talking_face``) on ``sys.path``;
  * stubs ``omegaconf.listconfig`` (lazy import at openaimodel.py:476; omegaconf is not installed);
  * provides ``LatentDiffusionStandIn`` — a 25-line ``nn.Module`` that reproduces the schedule
    buffers of ``DDPM.register_schedule`` (ddpm.py:117-142) with the reference's own
    ``make_beta_schedule`` and the ``apply_model``/``DiffusionWrapper`` semantics
    (ddpm.py:893-902,989,1404-1412; talking_face ddpm2cond.py:911-945,1307-1311);
  * provides ``CpuDDIMSampler`` — the reference ``DDIMSampler`` with ``register_buffer``
    overridden, because the stock one force-moves every buffer to ``cuda`` (ddim.py:18-22).
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch
import torch.nn as nn

REFERENCE_ROOT = os.environ.get("LDM_REFERENCE_ROOT", "/root/reference")


def available(flavour: str = "face_reenactment") -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, flavour, "ldm"))


def _purge_ldm_modules():
    for k in [k for k in sys.modules if k == "ldm" or k.startswith("ldm.")]:
        del sys.modules[k]


def import_reference(flavour: str = "face_reenactment"):
    """Returns the reference's ``ldm`` package for ``flavour`` ('face_reenactment' | 'talking_face')."""
    if not available(flavour):
        raise RuntimeError(f"reference not present under {REFERENCE_ROOT} (it never is on the GPU box)")
    root = os.path.join(REFERENCE_ROOT, flavour)
    if "omegaconf" not in sys.modules:
        oc = types.ModuleType("omegaconf")
        lc = types.ModuleType("omegaconf.listconfig")

        class ListConfig(list):
            pass

        lc.ListConfig = ListConfig
        oc.listconfig = lc
        oc.ListConfig = ListConfig
        sys.modules["omegaconf"] = oc
        sys.modules["omegaconf.listconfig"] = lc
    cur = sys.modules.get("ldm")
    if cur is not None and not getattr(cur, "__path__", [""])[0].startswith(root):
        _purge_ldm_modules()
    if root not in sys.path:
        sys.path.insert(0, root)
    else:  # make it the first hit
        sys.path.remove(root)
        sys.path.insert(0, root)
    import importlib
    return importlib.import_module("ldm")


def reference_unet(cfg: dict, flavour: str = "face_reenactment"):
    import_reference(flavour)
    from ldm.modules.diffusionmodules.openaimodel import UNetModel
    return UNetModel(**cfg)


class LatentDiffusionStandIn(nn.Module):
    """Minimal model handle a ``DDIMSampler`` needs: ``num_timesteps, betas, alphas_cumprod,
    alphas_cumprod_prev, device, apply_model``.  ``two_cond=True`` mirrors the
    talking-face ``apply_model(x, t, c12, c34)`` (concat c34 on channels, c12 as context)."""

    def __init__(self, unet: nn.Module, make_beta_schedule=None, linear_start=0.0015, linear_end=0.0205,
                 timesteps=1000, two_cond=False):
        super().__init__()
        if make_beta_schedule is None:
            from ldm.modules.diffusionmodules.util import make_beta_schedule
        betas = make_beta_schedule("linear", timesteps, linear_start=linear_start, linear_end=linear_end)
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
        self.unet = unet
        self.two_cond = two_cond

    @property
    def device(self):
        return self.betas.device

    def apply_model(self, x, t, c, c34=None):
        if self.two_cond:   # talking_face ddpm2cond.py:1307-1311
            return self.unet(torch.cat([x, c34], dim=1), t, context=c)
        return self.unet(x, t, context=c)  # ddpm.py:1410-1412, conditioning_key 'crossattn'

    def q_sample(self, x_start, t, noise=None):  # ddpm.py:273-276
        noise = torch.randn_like(x_start) if noise is None else noise
        e = lambda a: a.gather(-1, t).reshape(-1, 1, 1, 1)
        return e(self.sqrt_alphas_cumprod) * x_start + e(self.sqrt_one_minus_alphas_cumprod) * noise


def cpu_ddim_sampler(model, flavour: str = "face_reenactment"):
    """The reference DDIMSampler (ddim.py for FR, ddim2cond.py for TF) runnable without CUDA."""
    import_reference(flavour)
    if flavour == "face_reenactment":
        from ldm.models.diffusion.ddim import DDIMSampler
    else:
        from ldm.models.diffusion.ddim2cond import DDIMSampler

    class CpuDDIMSampler(DDIMSampler):
        def register_buffer(self, name, attr):  # ddim.py:18-22 without the hard-coded cuda move
            if type(attr) == torch.Tensor and attr.device != model.device:
                attr = attr.to(model.device)
            setattr(self, name, attr)

    return CpuDDIMSampler(model)


def cpu_manipulation_sampler(model):
    """The reference's DDIM-inversion sampler runnable here: ``ddim2.DDIMSampler2`` (FR ddim2.py:53) provides the
    strength-scaled schedule and ``p_sample_ddim``; ``q_sample_ddim`` and ``latent_manipulation`` exist only inside
    the private sampler class of FR/latent_manipulation.py (whose module-level imports — omegaconf, the AffectNet
    dataset — are absent here), so their UNMODIFIED source text is cut out of that file and bound to a subclass."""
    import re
    import textwrap
    import numpy as np
    from tqdm import tqdm
    import_reference("face_reenactment")
    from ldm.models.diffusion.ddim2 import DDIMSampler2
    src = open(os.path.join(REFERENCE_ROOT, "face_reenactment", "latent_manipulation.py")).read()

    def cut(name):
        m = re.search(r"^(    (?:@torch\.no_grad\(\)\n    )?def %s\(.*?)(?=^    (?:@|def )|^if __name__|\Z)" % name, src, re.S | re.M)
        assert m, name
        return textwrap.dedent(m.group(1))

    ns = {"torch": torch, "np": np, "tqdm": lambda it, **k: it}
    exec(cut("q_sample_ddim"), ns)
    exec(cut("latent_manipulation"), ns)

    class CpuManipulationSampler(DDIMSampler2):
        q_sample_ddim = ns["q_sample_ddim"]
        latent_manipulation = ns["latent_manipulation"]

        def register_buffer(self, name, attr):  # ddim2.py:60-64 without the hard-coded cuda move
            if type(attr) == torch.Tensor and attr.device != model.device:
                attr = attr.to(model.device)
            setattr(self, name, attr)

    return CpuManipulationSampler(model)


def cpu_progressive_sampler(model, record: list = None):
    """The talking-face progressive sampler runnable here: the private ``DDIMSampler`` class of
    TF/progressive_sampling_difftalk.py:55-319 (its own ``make_schedule``, ``p_sample_ddim`` :207-243 and the
    autoregressive ``progressive_sampling`` :246-319 with ``zid = img`` at :317).  The script's module-level imports
    (taming's MEAD dataset, omegaconf, PIL) are absent here, so the UNMODIFIED source text of the class is cut out of
    the file and executed against the reference's own ``ldm.modules.diffusionmodules.util`` helpers.  ``record``, when
    given, receives every 4-D ``torch.randn`` draw of the class body in call order (the per-frame x_T; the per-step
    ``noise_like`` draws happen inside the reference's util module) so that a GPU run can be fed the same starting noise."""
    import re
    from tqdm import tqdm  # noqa: F401
    import_reference("talking_face")
    from ldm.modules.diffusionmodules.util import (make_ddim_sampling_parameters, make_ddim_timesteps, noise_like,
                                                   extract_into_tensor)
    src = open(os.path.join(REFERENCE_ROOT, "talking_face", "progressive_sampling_difftalk.py")).read()
    m = re.search(r"^class DDIMSampler\(object\):.*?(?=^def main\(\))", src, re.S | re.M)
    assert m, "class DDIMSampler not found in progressive_sampling_difftalk.py"

    class _TorchProxy:
        """``torch`` with ``randn`` recorded (everything else is the real module)."""

        def __getattr__(self, name):
            return getattr(torch, name)

        @staticmethod
        def randn(*a, **k):
            out = torch.randn(*a, **k)
            if record is not None and out.dim() == 4:
                record.append(out.clone())
            return out

    ns = {"torch": _TorchProxy(), "np": np, "tqdm": lambda it, **k: it, "noise_like": noise_like,
          "make_ddim_sampling_parameters": make_ddim_sampling_parameters, "make_ddim_timesteps": make_ddim_timesteps,
          "extract_into_tensor": extract_into_tensor, "__name__": "progressive_sampling_difftalk_cut"}
    exec(m.group(0), ns)
    base = ns["DDIMSampler"]

    class CpuProgressiveSampler(base):
        def register_buffer(self, name, attr):  # :62-66 without the hard-coded cuda move
            if type(attr) == torch.Tensor and attr.device != model.device:
                attr = attr.to(model.device)
            setattr(self, name, attr)

    return CpuProgressiveSampler(model)
