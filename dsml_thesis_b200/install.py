"""``install()`` — the post-import patch that makes the reference pick up the B200 classes.

Every shipped YAML names ``target: ldm.modules.diffusionmodules.openaimodel.UNetModel`` and
``instantiate_from_config`` resolves it with ``importlib.import_module`` + ``getattr`` at call time
(ldm/util.py:78-93), so replacing the module attribute is enough for ``main.py`` and every sampling script.
The samplers are imported BY NAME into ``ldm.models.diffusion.ddpm`` (ddpm.py:26) and friends, so those
importing modules are patched too when they are already loaded.

Usage (no reference file edited):  put this repo on ``PYTHONPATH`` and either run
``python -c "import dsml_thesis_b200; dsml_thesis_b200.install()" ...`` / add the two lines to a
``sitecustomize.py``, or set ``LDMB200_AUTOINSTALL=1`` with the ``sitecustomize`` shipped in INTEGRATION.md.

``ldm.models.autoencoder.VQModelInterface.decode`` is wrapped so that CUDA latents decode through ``VQDecoder``
(built lazily per instance from the live module's own weights).

It also wraps ``ldm.modules.ema.LitEma.copy_to`` / ``restore`` (ema.py:46-76): they overwrite ``param.data`` in
place, which bumps neither ``_version`` nor ``data_ptr`` — the UNet's packed bf16 weights must be told.
"""
from __future__ import annotations

import importlib
import sys

_saved = {}


def _patch(modname: str, attr: str, value) -> bool:
    try:
        mod = importlib.import_module(modname)
    except Exception:
        return False
    if not hasattr(mod, attr):
        return False
    _saved.setdefault((modname, attr), getattr(mod, attr))
    setattr(mod, attr, value)
    return True


def install(verbose: bool = False):
    """Patches the reference's dotted targets.  Needs the reference's ``ldm`` package importable (it is when a
    reference script runs from its own directory).  Returns the list of patched ``module.attr`` names."""
    from .unet import UNetModel
    from .sampler import DDIMSampler, DDIMSampler2, DDIMSampler2Cond, DDIMSampler2CondFR
    done = []

    def p(mod, attr, val):
        if _patch(mod, attr, val):
            done.append(f"{mod}.{attr}")

    p("ldm.modules.diffusionmodules.openaimodel", "UNetModel", UNetModel)
    p("ldm.models.diffusion.ddim", "DDIMSampler", DDIMSampler)
    p("ldm.models.diffusion.ddim2", "DDIMSampler2", DDIMSampler2)   # strength-scaled grid + DDIM inversion (FR only)
    # the two ddim2cond flavours differ by their dict keys: talking_face uses 'class_label_&_audio'
    try:
        src = open(importlib.import_module("ldm.models.diffusion.ddim2cond").__file__).read()
        flavour = DDIMSampler2Cond if "class_label_&_audio" in src else DDIMSampler2CondFR
        p("ldm.models.diffusion.ddim2cond", "DDIMSampler", flavour)
    except Exception:
        flavour = None
    # modules that did `from ... import DDIMSampler` before install() ran
    for name, mod in list(sys.modules.items()):
        if not name.startswith("ldm.models.diffusion.") or mod is None:
            continue
        cur = getattr(mod, "DDIMSampler", None)
        if cur is None or cur in (DDIMSampler, DDIMSampler2Cond, DDIMSampler2CondFR):
            continue
        origin = getattr(cur, "__module__", "")
        if origin == "ldm.models.diffusion.ddim":
            p(name, "DDIMSampler", DDIMSampler)
        elif origin == "ldm.models.diffusion.ddim2cond" and flavour is not None:
            p(name, "DDIMSampler", flavour)
    # EMA weight swaps must invalidate the packed weights
    try:
        ema = importlib.import_module("ldm.modules.ema")
        lit = ema.LitEma
        if not getattr(lit, "_ldmb200_wrapped", False):
            def wrap(fn):
                def inner(self, *a, **k):
                    out = fn(self, *a, **k)
                    for obj in list(a) + list(k.values()):
                        mods = obj.modules() if hasattr(obj, "modules") else ()
                        for m in mods:
                            if isinstance(m, UNetModel):
                                m.invalidate_weight_cache()
                    return out
                return inner
            _saved[("ldm.modules.ema", "LitEma.copy_to")] = lit.copy_to
            _saved[("ldm.modules.ema", "LitEma.restore")] = lit.restore
            lit.copy_to = wrap(lit.copy_to)
            lit.restore = wrap(lit.restore)
            lit._ldmb200_wrapped = True
            done.append("ldm.modules.ema.LitEma.copy_to/restore")
    except Exception:
        pass
    # first-stage decode: VQModelInterface.decode -> VQDecoder, built lazily per instance
    try:
        ae = importlib.import_module("ldm.models.autoencoder")
        vqi = ae.VQModelInterface
        if not getattr(vqi, "_ldmb200_wrapped", False):
            from .vqdecode import accelerated_decode, accelerated_encode
            _saved[("ldm.models.autoencoder", "VQModelInterface.decode")] = vqi.decode
            _saved[("ldm.models.autoencoder", "VQModelInterface.encode")] = vqi.encode
            vqi.decode = accelerated_decode(vqi.decode)
            vqi.encode = accelerated_encode(vqi.encode)
            vqi._ldmb200_wrapped = True
            done.append("ldm.models.autoencoder.VQModelInterface.decode")
            done.append("ldm.models.autoencoder.VQModelInterface.encode")
    except Exception:
        pass
    # audio conditioning encoder (talking_face only): one fused launch per frame when no gradient is needed
    try:
        enc = importlib.import_module("ldm.modules.encoders.modules")
        cta = getattr(enc, "Conv1DTemporalAttention", None)
        if cta is not None and not getattr(cta, "_ldmb200_wrapped", False):
            from .cond import accelerated_forward
            _saved[("ldm.modules.encoders.modules", "Conv1DTemporalAttention.forward")] = cta.forward
            cta.forward = accelerated_forward(cta.forward)
            cta._ldmb200_wrapped = True
            done.append("ldm.modules.encoders.modules.Conv1DTemporalAttention.forward")
    except Exception:
        pass
    if verbose:
        print("dsml_thesis_b200.install():", ", ".join(done) or "nothing patched (is `ldm` importable?)")
    return done


def uninstall():
    for (modname, attr), val in list(_saved.items()):
        mod = sys.modules.get(modname)
        if mod is None:
            continue
        if "." in attr:
            cls, meth = attr.split(".")
            setattr(getattr(mod, cls), meth, val)
            getattr(mod, cls)._ldmb200_wrapped = False
        else:
            setattr(mod, attr, val)
    _saved.clear()
