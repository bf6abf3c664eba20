"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol include/ldmb200.h declares,
the drop-in module reproduces the reference's parameter inventory and schedule arithmetic, install() patches the
reference's dotted targets, and the product path fails loudly without a GPU."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

import dsml_thesis_b200 as P
from dsml_thesis_b200 import _cabi
from oracle import ldm_oracle as O
from oracle import ref_shim as R

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "ldmb200.h")).read()
    declared = set(re.findall(r"\b(ldmb200_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"ldmb200_gemm_desc", "ldmb200_pack_job"}
    assert declared == set(_cabi.SIGNATURES), declared ^ set(_cabi.SIGNATURES)
    lib = P.load_library()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ldmb200_abi_version() == 5


def test_struct_layouts_match_the_header():
    # field order / count of the ctypes mirrors vs the C structs (names are identical by construction)
    hdr = open(os.path.join(ROOT, "include", "ldmb200.h")).read()
    for cname, cls in (("ldmb200_gemm_desc", _cabi.GemmDesc), ("ldmb200_pack_job", _cabi.PackJob),
                       ("ldmb200_wgrad_desc", _cabi.WgradDesc), ("ldmb200_linear_bwd_job", _cabi.LinearBwdJob)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), hdr, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            parts = decl.split(",")
            first = parts[0].split()[-1].lstrip("*")
            names.append(first)
            names += [p.strip().lstrip("*") for p in parts[1:]]
        assert names == [f[0].rstrip("_") if f[0] == "in_" else f[0] for f in cls._fields_], (cname, names)


@pytest.mark.parametrize("cfg", [O.TINY_UNET, O.FR_UNET, O.TF_UNET])
def test_parameter_inventory_matches_reference(cfg):
    m = P.UNetModel(**cfg)
    sd = m.state_dict()
    shapes = O.param_shapes(cfg)
    assert set(sd) == set(shapes)
    assert all(tuple(sd[k].shape) == shapes[k] and sd[k].dtype == torch.float32 for k in sd)
    assert len(list(m.buffers())) == 0
    # the 34 zero_module layers of the reference are zero here too
    zero = [k for k, v in sd.items() if k.endswith(".weight") and v.dim() > 1 and float(v.abs().max()) == 0.0]
    if cfg is not O.TINY_UNET:
        assert len(zero) == 34


@pytest.mark.skipif(not R.available(), reason="reference tree only exists in the build container")
def test_state_dict_loads_into_the_real_reference_and_back():
    cfg = O.TINY_UNET
    ref = R.reference_unet(cfg)
    m = P.UNetModel(**cfg)
    m.load_state_dict(ref.state_dict(), strict=True)
    ref.load_state_dict(m.state_dict(), strict=True)


def test_unsupported_configurations_fail_at_construction():
    for kw in (dict(use_scale_shift_norm=True), dict(resblock_updown=True), dict(dims=3), dict(dropout=0.1),
               dict(n_embed=10)):
        with pytest.raises(NotImplementedError):
            P.UNetModel(**dict(O.TINY_UNET, **kw))


def test_sampler_schedule_matches_reference_golden(golden_schedule):
    h = P.DiffusionHandle(torch.nn.Identity())
    for k in ("betas", "alphas_cumprod", "alphas_cumprod_prev"):
        assert np.array_equal(getattr(h, k).numpy(), golden_schedule["ddpm/" + k])
    for S in (50, 200, 7):
        for eta in (0.0, 1.0):
            s = P.DDIMSampler(h)
            s.make_schedule(S, ddim_eta=eta, verbose=False)
            key = f"ddim/S{S}_eta{int(eta)}"
            assert np.array_equal(s.ddim_timesteps, golden_schedule[key + "/timesteps"])
            for f, kk in (("a_t", "a_t"), ("a_prev", "a_prev"), ("sigma_t", "sigma"), ("sqrt_one_minus_at", "s1")):
                assert np.array_equal(np.asarray(s._coef[kk], dtype=np.float32), golden_schedule[f"{key}/{f}"])


def test_no_cpu_fallback():
    m = P.UNetModel(**O.TINY_UNET).eval()
    x = torch.randn(1, 3, 8, 8)
    with pytest.raises(P.Ldmb200Error):
        with torch.no_grad():
            m(x, torch.tensor([5]), context=torch.randn(1, 1, 64))
    h = P.DiffusionHandle(m)
    s = P.DDIMSampler(h)
    s.make_schedule(5, verbose=False)
    with pytest.raises(P.Ldmb200Error):
        s.p_sample_ddim(x, torch.randn(1, 1, 64), torch.tensor([5]), index=0)


def test_missing_library_is_a_loud_error():
    code = ("import os, sys; sys.path.insert(0, %r); os.environ['LDMB200_LIB'] = '/nonexistent/lib.so';"
            "import dsml_thesis_b200 as P\n"
            "try:\n    P.load_library()\nexcept P.Ldmb200Error as e:\n    print('LOUD', e)\n") % ROOT
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert "LOUD" in out.stdout and "no CPU/PyTorch fallback" in out.stdout, out.stdout + out.stderr


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "dsml_thesis_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


@pytest.mark.skipif(not R.available(), reason="reference tree only exists in the build container")
def test_install_patches_the_reference_targets():
    R.import_reference("face_reenactment")
    import ldm.models.diffusion.ddim as ddim_mod
    import ldm.modules.diffusionmodules.openaimodel as oam
    ref_unet, ref_smp = oam.UNetModel, ddim_mod.DDIMSampler
    try:
        done = P.install()
        assert "ldm.modules.diffusionmodules.openaimodel.UNetModel" in done
        assert oam.UNetModel is P.UNetModel and ddim_mod.DDIMSampler is P.DDIMSampler
        # instantiate_from_config semantics (ldm/util.py:78-93): importlib + getattr at call time
        import importlib
        cls = getattr(importlib.import_module("ldm.modules.diffusionmodules.openaimodel"), "UNetModel")
        assert isinstance(cls(**O.TINY_UNET), P.UNetModel)
    finally:
        P.uninstall()
    assert oam.UNetModel is ref_unet and ddim_mod.DDIMSampler is ref_smp


# ------------------------------------------------------------------------------------------ first-stage decode (8f n2)
def test_vq_decoder_parameter_inventory_matches_the_reference_names():
    """VQDecoder's state dict = the decode-side keys of the reference VQModel (oracle.vq_oracle.vq_param_shapes, proven
    against the real modules by load_state_dict(strict=True) in tests/golden/make_golden.py)."""
    import dsml_thesis_b200 as P
    from oracle import vq_oracle as V
    for cfg in (V.VQ_TINY, V.VQ_F4):
        m = P.VQDecoder(**cfg, lossconfig={"target": "torch.nn.Identity"}, ckpt_path=None)
        mine = {k: tuple(v.shape) for k, v in m.state_dict().items()}
        assert mine == V.vq_param_shapes(cfg)
        assert not any(p.requires_grad for p in m.parameters())
        enc = P.VQEncoder(**cfg)
        assert {k: tuple(v.shape) for k, v in enc.state_dict().items()} == V.vq_encoder_param_shapes(cfg)
    # a full first-stage checkpoint (encoder / quant_conv / loss keys too) loads with strict=False and nothing missing
    sd = dict(V.init_vq_state_dict(V.VQ_TINY, 0))
    sd["encoder.conv_in.weight"] = torch.zeros(32, 3, 3, 3)
    sd["quant_conv.weight"] = torch.zeros(3, 3, 1, 1)
    missing, unexpected = P.VQDecoder(**V.VQ_TINY).load_state_dict(sd, strict=False)
    assert not missing and sorted(unexpected) == ["encoder.conv_in.weight", "quant_conv.weight"]


def test_vq_decoder_has_no_cpu_fallback_and_rejects_other_variants():
    import dsml_thesis_b200 as P
    from oracle import vq_oracle as V
    m = P.VQDecoder(**V.VQ_TINY)
    with pytest.raises(P.Ldmb200Error):
        m.decode(torch.zeros(1, 3, 8, 8))
    dd = dict(V.VQ_TINY["ddconfig"], attn_type="linear")
    with pytest.raises(NotImplementedError):
        P.VQDecoder(dd, 3, 512)


def test_install_wraps_first_stage_decode(monkeypatch):
    """``ldm.models.autoencoder`` needs pytorch_lightning (absent here), so a stand-in module with the reference's class
    name is registered: install() must wrap ``VQModelInterface.decode`` at class level, keep CPU calls on the reference
    method, and uninstall() must restore it."""
    import types
    import dsml_thesis_b200 as P
    R.import_reference("face_reenactment")
    calls = []

    class VQModelInterface(torch.nn.Module):
        def decode(self, h, force_not_quantize=False):
            calls.append(force_not_quantize)
            return h

        def encode(self, x):
            calls.append("encode")
            return x

    mod = types.ModuleType("ldm.models.autoencoder")
    mod.VQModelInterface = VQModelInterface
    monkeypatch.setitem(sys.modules, "ldm.models.autoencoder", mod)
    ref_decode, ref_encode = VQModelInterface.decode, VQModelInterface.encode
    try:
        done = P.install()
        assert "ldm.models.autoencoder.VQModelInterface.decode" in done
        assert VQModelInterface.decode is not ref_decode
        x = torch.zeros(1, 3, 4, 4)
        assert VQModelInterface().decode(x, True) is x and calls == [True]     # CPU tensor: the reference's own method
        assert "ldm.models.autoencoder.VQModelInterface.encode" in done
        assert VQModelInterface().encode(x) is x and calls == [True, "encode"]
    finally:
        P.uninstall()
    assert VQModelInterface.decode is ref_decode and VQModelInterface.encode is ref_encode
