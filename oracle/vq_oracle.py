"""CPU oracle of the first stage: decode and encode (row n4) — TEST INFRASTRUCTURE ONLY.

Functional fp32 restatement (torch CPU ops over a plain state dict) of ``VQModelInterface.decode``
(ldm/models/autoencoder.py:274-283): ``VectorQuantizer2.forward`` (taming/modules/vqvae/quantize.py:285-312) →
``post_quant_conv`` (autoencoder.py:43) → ``Decoder.forward`` (ldm/modules/diffusionmodules/model.py:535-569) with
``ResnetBlock`` (model.py:120-142, temb=None), ``AttnBlock`` (172-198), ``Upsample`` (53-57), ``Normalize`` (38-39,
GroupNorm 32 groups eps 1e-6) and the swish ``nonlinearity`` (33-35).

The encode half restates ``VQModelInterface.encode`` (autoencoder.py:269-272): ``Encoder.forward`` (model.py:433-459) with
the (0,1,0,1)-padded stride-2 ``Downsample`` (72-76) → ``quant_conv``.

Pinned by ``tests/test_oracle_golden.py`` against outputs of the REAL reference modules
(``tests/golden/make_golden.py: vq_goldens / vq_encode_goldens`` → ``tests/golden/vqdecode.npz``, ``vqencode.npz``).  Only ``tests/``, ``smoke()`` and
``bench.py``'s CPU-baseline legs may import this module; the product never does.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# first_stage_config of every shipped LDM config (FR affectnet-128-ldm-vq-f4.yaml:43-65, TF mead-128-ldm-f4.yaml:44-66)
VQ_F4 = dict(embed_dim=3, n_embed=16384,
             ddconfig=dict(double_z=False, z_channels=3, resolution=128, in_channels=3, out_ch=3, ch=128,
                           ch_mult=[1, 2, 4], num_res_blocks=2, attn_resolutions=[32], dropout=0.0))
# same topology (3 levels, attention at the lowest level and in the middle, nin_shortcut at level changes), small enough
# for CPU tests
VQ_TINY = dict(embed_dim=3, n_embed=512,
               ddconfig=dict(double_z=False, z_channels=3, resolution=32, in_channels=3, out_ch=3, ch=32,
                             ch_mult=[1, 2, 4], num_res_blocks=1, attn_resolutions=[8], dropout=0.0))


def decoder_structure(dd: dict):
    """Walks ``Decoder.__init__`` (model.py:478-533): returns (block_in at the lowest level, per-level list of
    (in_ch, out_ch, has_attn) in FORWARD order i.e. lowest resolution first, and which levels upsample)."""
    ch, ch_mult, nrb = dd["ch"], list(dd["ch_mult"]), dd["num_res_blocks"]
    nres = len(ch_mult)
    block_in = ch * ch_mult[nres - 1]
    curr_res = dd["resolution"] // 2 ** (nres - 1)
    first = block_in
    levels = {}
    for i_level in reversed(range(nres)):
        block_out = ch * ch_mult[i_level]
        blocks = []
        for _ in range(nrb + 1):
            blocks.append((block_in, block_out, curr_res in dd["attn_resolutions"]))
            block_in = block_out
        levels[i_level] = dict(blocks=blocks, upsample=(i_level != 0), ch=block_in)
        if i_level != 0:
            curr_res *= 2
    return first, levels, block_in


def vq_param_shapes(cfg: dict) -> Dict[str, Tuple[int, ...]]:
    """Names and shapes of the ``VQModelInterface`` state-dict entries the decode path reads."""
    dd = cfg["ddconfig"]
    e, zc = cfg["embed_dim"], dd["z_channels"]
    sh: Dict[str, Tuple[int, ...]] = {}
    sh["quantize.embedding.weight"] = (cfg["n_embed"], e)
    sh["post_quant_conv.weight"] = (zc, e, 1, 1)
    sh["post_quant_conv.bias"] = (zc,)

    def conv(p, cin, cout, k):
        sh[p + ".weight"] = (cout, cin, k, k)
        sh[p + ".bias"] = (cout,)

    def norm(p, c):
        sh[p + ".weight"] = (c,)
        sh[p + ".bias"] = (c,)

    def resnet(p, cin, cout):
        norm(p + ".norm1", cin)
        conv(p + ".conv1", cin, cout, 3)
        norm(p + ".norm2", cout)
        conv(p + ".conv2", cout, cout, 3)
        if cin != cout:
            conv(p + ".nin_shortcut", cin, cout, 1)

    def attn(p, c):
        norm(p + ".norm", c)
        for n in ("q", "k", "v", "proj_out"):
            conv(f"{p}.{n}", c, c, 1)

    first, levels, last = decoder_structure(dd)
    conv("decoder.conv_in", zc, first, 3)
    resnet("decoder.mid.block_1", first, first)
    attn("decoder.mid.attn_1", first)
    resnet("decoder.mid.block_2", first, first)
    for i_level, lv in levels.items():
        for j, (cin, cout, has_attn) in enumerate(lv["blocks"]):
            resnet(f"decoder.up.{i_level}.block.{j}", cin, cout)
            if has_attn:
                attn(f"decoder.up.{i_level}.attn.{j}", cout)
        if lv["upsample"]:
            conv(f"decoder.up.{i_level}.upsample.conv", lv["ch"], lv["ch"], 3)
    norm("decoder.norm_out", last)
    conv("decoder.conv_out", last, dd["out_ch"], 3)
    return sh


def init_vq_state_dict(cfg: dict, seed: int = 0) -> Dict[str, Tensor]:
    """Deterministic synthetic first-stage weights.  Convs ~ sqrt(3)·U(-1,1)/sqrt(fan_in), biases 0.05·N, norm gains
    1+0.1·N; the codebook is N(0,1) like a trained one (the freshly initialised reference codebook,
    U(-1/n_e, 1/n_e), makes all distances collide within fp32 rounding of |z|² — a degenerate argmin)."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    for name, shape in sorted(vq_param_shapes(cfg).items()):
        if name == "quantize.embedding.weight":
            sd[name] = torch.randn(shape, generator=g)
        elif len(shape) >= 2:
            fan_in = int(np.prod(shape[1:]))
            sd[name] = ((torch.rand(shape, generator=g) * 2 - 1) * math.sqrt(3.0 / fan_in)).contiguous()
        elif "norm" in name:
            sd[name] = (1.0 + 0.1 * torch.randn(shape, generator=g)) if name.endswith("weight") \
                else 0.1 * torch.randn(shape, generator=g)
        else:
            sd[name] = 0.05 * torch.randn(shape, generator=g)
    return sd


def swish(x: Tensor) -> Tensor:
    """model.py:33-35."""
    return x * torch.sigmoid(x)


def normalize(sd, p: str, x: Tensor) -> Tensor:
    """model.py:38-39: GroupNorm(32 groups, eps 1e-6, affine)."""
    return F.group_norm(x, 32, sd[p + ".weight"], sd[p + ".bias"], eps=1e-6)


def resnet_block(sd, p: str, x: Tensor) -> Tensor:
    """model.py:120-142 with temb=None, dropout 0."""
    h = F.conv2d(swish(normalize(sd, p + ".norm1", x)), sd[p + ".conv1.weight"], sd[p + ".conv1.bias"], padding=1)
    h = F.conv2d(swish(normalize(sd, p + ".norm2", h)), sd[p + ".conv2.weight"], sd[p + ".conv2.bias"], padding=1)
    if p + ".nin_shortcut.weight" in sd:
        x = F.conv2d(x, sd[p + ".nin_shortcut.weight"], sd[p + ".nin_shortcut.bias"])
    return x + h


def attn_block(sd, p: str, x: Tensor) -> Tensor:
    """model.py:172-198: single-head attention over the h*w positions, d = C."""
    h_ = normalize(sd, p + ".norm", x)
    q = F.conv2d(h_, sd[p + ".q.weight"], sd[p + ".q.bias"])
    k = F.conv2d(h_, sd[p + ".k.weight"], sd[p + ".k.bias"])
    v = F.conv2d(h_, sd[p + ".v.weight"], sd[p + ".v.bias"])
    b, c, hh, ww = q.shape
    q = q.reshape(b, c, hh * ww).permute(0, 2, 1)
    k = k.reshape(b, c, hh * ww)
    w_ = torch.bmm(q, k) * (int(c) ** (-0.5))
    w_ = F.softmax(w_, dim=2)
    v = v.reshape(b, c, hh * ww)
    h_ = torch.bmm(v, w_.permute(0, 2, 1)).reshape(b, c, hh, ww)
    h_ = F.conv2d(h_, sd[p + ".proj_out.weight"], sd[p + ".proj_out.bias"])
    return x + h_


def vector_quantize(sd, z: Tensor) -> Tuple[Tensor, Tensor]:
    """quantize.py:285-312 (inference outputs only): returns (z_q NCHW, indices [B*H*W])."""
    e = sd["quantize.embedding.weight"]
    zp = z.permute(0, 2, 3, 1).contiguous()
    zf = zp.view(-1, e.shape[1])
    d = torch.sum(zf ** 2, dim=1, keepdim=True) + torch.sum(e ** 2, dim=1) - 2 * torch.einsum("bd,dn->bn", zf, e.t())
    idx = torch.argmin(d, dim=1)
    z_q = e[idx].view(zp.shape)
    z_q = zp + (z_q - zp)
    return z_q.permute(0, 3, 1, 2).contiguous(), idx


def decoder_forward(sd, cfg: dict, z: Tensor) -> Tensor:
    """model.py:535-569 (give_pre_end=False, tanh_out=False)."""
    dd = cfg["ddconfig"]
    _, levels, _ = decoder_structure(dd)
    h = F.conv2d(z, sd["decoder.conv_in.weight"], sd["decoder.conv_in.bias"], padding=1)
    h = resnet_block(sd, "decoder.mid.block_1", h)
    h = attn_block(sd, "decoder.mid.attn_1", h)
    h = resnet_block(sd, "decoder.mid.block_2", h)
    for i_level, lv in levels.items():          # lowest resolution first
        for j, (_, _, has_attn) in enumerate(lv["blocks"]):
            h = resnet_block(sd, f"decoder.up.{i_level}.block.{j}", h)
            if has_attn:
                h = attn_block(sd, f"decoder.up.{i_level}.attn.{j}", h)
        if lv["upsample"]:
            h = F.interpolate(h, scale_factor=2.0, mode="nearest")
            p = f"decoder.up.{i_level}.upsample.conv"
            h = F.conv2d(h, sd[p + ".weight"], sd[p + ".bias"], padding=1)
    h = swish(normalize(sd, "decoder.norm_out", h))
    return F.conv2d(h, sd["decoder.conv_out.weight"], sd["decoder.conv_out.bias"], padding=1)


@torch.no_grad()
def vq_decode(sd, cfg: dict, h: Tensor, force_not_quantize: bool = False):
    """autoencoder.py:274-283.  Returns (image, indices or None)."""
    idx = None
    quant = h
    if not force_not_quantize:
        quant, idx = vector_quantize(sd, h)
    quant = F.conv2d(quant, sd["post_quant_conv.weight"], sd["post_quant_conv.bias"])
    return decoder_forward(sd, cfg, quant), idx


# ----------------------------------------------------------------------------------------
# encode half (encode_first_stage per frame) — VQModelInterface.encode, autoencoder.py:269-272
# ----------------------------------------------------------------------------------------

def encoder_structure(dd: dict):
    """Walks ``Encoder.__init__`` (model.py:383-412): per-level list of (in_ch, out_ch, has_attn), which levels
    downsample, and the channel count at the lowest level."""
    ch, ch_mult, nrb = dd["ch"], list(dd["ch_mult"]), dd["num_res_blocks"]
    in_ch_mult = [1] + ch_mult
    curr_res = dd["resolution"]
    levels = []
    block_in = ch
    for i_level in range(len(ch_mult)):
        block_in, block_out = ch * in_ch_mult[i_level], ch * ch_mult[i_level]
        blocks = []
        for _ in range(nrb):
            blocks.append((block_in, block_out, curr_res in dd["attn_resolutions"]))
            block_in = block_out
        down = i_level != len(ch_mult) - 1
        levels.append(dict(blocks=blocks, downsample=down, ch=block_in))
        if down:
            curr_res //= 2
    return levels, block_in


def vq_encoder_param_shapes(cfg: dict) -> Dict[str, Tuple[int, ...]]:
    """Names and shapes of the ``VQModelInterface`` state-dict entries the encode path reads."""
    dd = cfg["ddconfig"]
    sh: Dict[str, Tuple[int, ...]] = {}

    def conv(p, cin, cout, k):
        sh[p + ".weight"] = (cout, cin, k, k)
        sh[p + ".bias"] = (cout,)

    def norm(p, c):
        sh[p + ".weight"] = (c,)
        sh[p + ".bias"] = (c,)

    def resnet(p, cin, cout):
        norm(p + ".norm1", cin)
        conv(p + ".conv1", cin, cout, 3)
        norm(p + ".norm2", cout)
        conv(p + ".conv2", cout, cout, 3)
        if cin != cout:
            conv(p + ".nin_shortcut", cin, cout, 1)

    def attn(p, c):
        norm(p + ".norm", c)
        for n in ("q", "k", "v", "proj_out"):
            conv(f"{p}.{n}", c, c, 1)

    levels, last = encoder_structure(dd)
    conv("encoder.conv_in", dd["in_channels"], dd["ch"], 3)
    for i_level, lv in enumerate(levels):
        for j, (cin, cout, has_attn) in enumerate(lv["blocks"]):
            resnet(f"encoder.down.{i_level}.block.{j}", cin, cout)
            if has_attn:
                attn(f"encoder.down.{i_level}.attn.{j}", cout)
        if lv["downsample"]:
            conv(f"encoder.down.{i_level}.downsample.conv", lv["ch"], lv["ch"], 3)
    resnet("encoder.mid.block_1", last, last)
    attn("encoder.mid.attn_1", last)
    resnet("encoder.mid.block_2", last, last)
    norm("encoder.norm_out", last)
    zc = 2 * dd["z_channels"] if dd.get("double_z", True) else dd["z_channels"]
    conv("encoder.conv_out", last, zc, 3)
    conv("quant_conv", zc, cfg["embed_dim"], 1)
    return sh


def init_vq_encoder_state_dict(cfg: dict, seed: int = 0) -> Dict[str, Tensor]:
    g = torch.Generator().manual_seed(seed + 1000)
    sd: Dict[str, Tensor] = {}
    for name, shape in sorted(vq_encoder_param_shapes(cfg).items()):
        if len(shape) >= 2:
            fan_in = int(np.prod(shape[1:]))
            sd[name] = ((torch.rand(shape, generator=g) * 2 - 1) * math.sqrt(3.0 / fan_in)).contiguous()
        elif "norm" in name:
            sd[name] = (1.0 + 0.1 * torch.randn(shape, generator=g)) if name.endswith("weight") \
                else 0.1 * torch.randn(shape, generator=g)
        else:
            sd[name] = 0.05 * torch.randn(shape, generator=g)
    return sd


def downsample(sd, p: str, x: Tensor) -> Tensor:
    """model.py:72-76: pad (0,1,0,1) with zeros, conv 3x3 stride 2 padding 0."""
    x = F.pad(x, (0, 1, 0, 1), mode="constant", value=0)
    return F.conv2d(x, sd[p + ".conv.weight"], sd[p + ".conv.bias"], stride=2, padding=0)


def encoder_forward(sd, cfg: dict, x: Tensor) -> Tensor:
    """model.py:433-459."""
    levels, _ = encoder_structure(cfg["ddconfig"])
    h = F.conv2d(x, sd["encoder.conv_in.weight"], sd["encoder.conv_in.bias"], padding=1)
    for i_level, lv in enumerate(levels):
        for j, (_, _, has_attn) in enumerate(lv["blocks"]):
            h = resnet_block(sd, f"encoder.down.{i_level}.block.{j}", h)
            if has_attn:
                h = attn_block(sd, f"encoder.down.{i_level}.attn.{j}", h)
        if lv["downsample"]:
            h = downsample(sd, f"encoder.down.{i_level}.downsample", h)
    h = resnet_block(sd, "encoder.mid.block_1", h)
    h = attn_block(sd, "encoder.mid.attn_1", h)
    h = resnet_block(sd, "encoder.mid.block_2", h)
    h = swish(normalize(sd, "encoder.norm_out", h))
    return F.conv2d(h, sd["encoder.conv_out.weight"], sd["encoder.conv_out.bias"], padding=1)


@torch.no_grad()
def vq_encode(sd, cfg: dict, x: Tensor) -> Tensor:
    """autoencoder.py:269-272: quant_conv(encoder(x)), no quantisation."""
    return F.conv2d(encoder_forward(sd, cfg, x), sd["quant_conv.weight"], sd["quant_conv.bias"])
