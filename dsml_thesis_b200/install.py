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
import importlib.util
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


def _exclude_unet_from_ddp():
    """``DistributedDataParallel`` must not manage the UNet's parameters: their gradients are written by the B200
    backward into a flat arena (no autograd accumulation hooks fire), so DDP's reducer would wait for them forever
    ("Expected to have finished reduction in the prior iteration") or, with ``find_unused_parameters``, skip them and let
    the replicas diverge.  The reference's ``main.py`` wraps the whole ``LatentDiffusion`` in Lightning DDP
    (main.py:531-543); this wrapper adds every parameter of a ``dsml_thesis_b200.UNetModel`` inside the wrapped module to
    DDP's documented ignore list before the stock constructor runs.  Their gradient exchange is
    ``dsml_thesis_b200.train.GradAllReduce``, attached automatically by the first multi-rank training forward."""
    import torch
    from .unet import UNetModel
    ddp = torch.nn.parallel.DistributedDataParallel
    if getattr(ddp, "_ldmb200_wrapped", False):
        return False
    orig = ddp.__init__

    def init(self, module, *a, **k):
        names = [f"{mn}.{pn}" if mn else pn for mn, m in module.named_modules() if isinstance(m, UNetModel)
                 for pn, _ in m.named_parameters(recurse=True)]
        if names:
            prev = list(getattr(module, "_ddp_params_and_buffers_to_ignore", []))
            ddp._set_params_and_buffers_to_ignore_for_model(module, prev + [n for n in names if n not in prev])
        return orig(self, module, *a, **k)

    _saved[("torch.nn.parallel", "DistributedDataParallel.__init__")] = orig
    ddp.__init__ = init
    ddp._ldmb200_wrapped = True
    return True


def install(verbose: bool = False):
    """Patches the reference's dotted targets.  Needs the reference's ``ldm`` package importable (it is when a
    reference script runs from its own directory).  Returns the list of patched ``module.attr`` names; targets that exist
    but could not be patched are reported with a ``RuntimeWarning`` (and ``UNetModel`` itself failing raises): a partially
    patched reference must not silently run its own slow path."""
    import warnings
    from .unet import UNetModel
    from .sampler import DDIMSampler, DDIMSampler2, DDIMSampler2Cond, DDIMSampler2CondFR
    done = []
    failed = []

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
    except ImportError:
        flavour = None
    except Exception as e:   # the module exists but could not be read / patched
        flavour = None
        failed.append(f"ldm.models.diffusion.ddim2cond.DDIMSampler ({e!r})")
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
    except ImportError:
        pass
    except Exception as e:
        failed.append(f"ldm.modules.ema.LitEma ({e!r})")
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
    except ImportError:       # needs pytorch_lightning / taming: absent where only the UNet path is exercised
        pass
    except Exception as e:
        failed.append(f"ldm.models.autoencoder.VQModelInterface ({e!r})")
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
    except ImportError:       # needs clip / kornia at module level
        pass
    except Exception as e:
        failed.append(f"ldm.modules.encoders.modules.Conv1DTemporalAttention ({e!r})")
    # multi-GPU training through the reference's own trainer (Lightning DDP, main.py:531-543)
    try:
        if _exclude_unet_from_ddp():
            done.append("torch.nn.parallel.DistributedDataParallel.__init__ (UNet parameters excluded; GradAllReduce syncs them)")
    except Exception as e:
        failed.append(f"torch.nn.parallel.DistributedDataParallel ({e!r})")
    ldm_present = importlib.util.find_spec("ldm") is not None
    if ldm_present and "ldm.modules.diffusionmodules.openaimodel.UNetModel" not in done:
        raise RuntimeError("dsml_thesis_b200.install(): `ldm` is importable but "
                           "ldm.modules.diffusionmodules.openaimodel.UNetModel could not be patched — the reference would "
                           "silently run its own UNet")
    if failed:
        warnings.warn("dsml_thesis_b200.install(): not patched (the reference's own implementation stays in place): "
                      + "; ".join(failed), RuntimeWarning)
    if verbose:
        print("dsml_thesis_b200.install():", ", ".join(done) or "nothing patched (is `ldm` importable?)")
    return done


def uninstall():
    for (modname, attr), val in list(_saved.items()):
        if modname == "torch.nn.parallel":
            import torch
            torch.nn.parallel.DistributedDataParallel.__init__ = val
            torch.nn.parallel.DistributedDataParallel._ldmb200_wrapped = False
            continue
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
