"""B200-native (sm_100a) latent-diffusion denoising path of GiannisPikoulis/dsml-thesis.

Public surface = the reference's own plugin interface for this path:

  * ``UNetModel``       — drop-in for ``ldm.modules.diffusionmodules.openaimodel.UNetModel``
  * ``DDIMSampler``     — drop-in for ``ldm.models.diffusion.ddim.DDIMSampler``
  * ``DDIMSampler2``    — drop-in for ``ldm.models.diffusion.ddim2.DDIMSampler2`` (strength-scaled grid) plus the DDIM
                          inversion ``q_sample_ddim`` / ``latent_manipulation`` of FR/latent_manipulation.py
  * ``DDIMSampler2Cond`` / ``DDIMSampler2CondFR`` — drop-ins for the two ``ddim2cond.DDIMSampler`` flavours
  * ``install()``       — patches those dotted targets after ``import ldm`` so ``main.py``, the YAML configs and
                          the sampling scripts pick the new classes up with no file edited
  * ``ddim_step``       — the fused x_{t-1} update as a function
  * ``VQDecoder`` / ``accelerate_first_stage`` — the decode half of ``ldm.models.autoencoder.VQModelInterface``
                          (nearest-code lookup, post_quant_conv, Decoder) on the same kernels
  * ``sample_sharded``  — ``DDIMSampler.sample`` for a global batch split over the GPUs of one box (one process
                          per GPU, no data-path collective, final all-gather)
  * ``Conv1DTemporalAttention`` — the talking-face audio conditioning encoder as one fused launch

All compute goes through ``libldmb200.so`` (C-ABI in ``include/ldmb200.h``); there is no CPU or PyTorch
fallback — importing works anywhere, running needs the built library and a B200.
"""
from ._cabi import Ldmb200Error, load as load_library  # noqa: F401
from .unet import UNetModel  # noqa: F401
from .sampler import DDIMSampler, DDIMSampler2, DDIMSampler2Cond, DDIMSampler2CondFR, ddim_step  # noqa: F401
from .handle import DiffusionHandle  # noqa: F401
from .install import install, uninstall  # noqa: F401
from .vqdecode import VQDecoder, VQEncoder, accelerate_first_stage  # noqa: F401
from .cond import Conv1DTemporalAttention  # noqa: F401
from .sharding import sample_sharded, shard_range, gather_shards, initial_noise  # noqa: F401

__all__ = ["UNetModel", "DDIMSampler", "DDIMSampler2", "DDIMSampler2Cond", "DDIMSampler2CondFR", "ddim_step", "install",
           "uninstall", "DiffusionHandle", "load_library", "Ldmb200Error", "sample_sharded", "shard_range",
           "gather_shards", "initial_noise", "VQDecoder", "VQEncoder", "accelerate_first_stage",
           "Conv1DTemporalAttention"]
