"""First stage on the B200 kernels.  Decode: drop-in for ``VQModelInterface.decode``
(ldm/models/autoencoder.py:274-283) = ``VectorQuantizer2`` nearest-code lookup (taming/modules/vqvae/quantize.py:285-312)
→ ``post_quant_conv`` → ``Decoder`` (ldm/modules/diffusionmodules/model.py:462-569).

``VQDecoder`` holds the same parameters under the same names as the reference ``VQModel`` (``quantize.embedding.weight``,
``post_quant_conv.*``, ``decoder.*``), so ``load_state_dict(first_stage_model.state_dict(), strict=False)`` picks up a
trained first stage (the encoder / loss entries are simply not ours).  ``decode`` compiles, per (batch, H, W, precision,
quantize), a static launch list over the same channels-last arena, weight pack and kernels as the UNet:

* lookup + post_quant_conv + channels-last staging: one launch (``ldmb200_vq_lookup_postconv``);
* every ``Conv2d`` 3x3 / 1x1: the tcgen05 implicit GEMM with fused bias / residual / GroupNorm-statistics epilogue;
* ``Normalize`` + swish: the fused GroupNorm(+SiLU) pass (eps 1e-6);
* ``AttnBlock`` (single head, d = C = 512, so not the d_head-32 flash kernel): q, k as 1x1 GEMMs; per sample
  S = q kᵀ and O = P v run through the SAME implicit-GEMM kernel with the sample's k and vᵀ as the "weight" operand
  (vᵀ = W_v · hᵀ is produced directly in that layout by swapping the operands of the v projection, and v's bias moves
  to the O epilogue because softmax rows sum to one); softmax rows by ``ldmb200_softmax_rows``.

Encode (row n4, ``encode_first_stage`` of the masked frame in progressive sampling): ``VQEncoder`` =
``quant_conv(Encoder(x))`` (autoencoder.py:269-272, model.py:368-459) as one more launch plan; its ``Downsample`` pads
(0,1,0,1) and strides by 2, which the implicit GEMM takes as ``pad_mode = 1``.

No CPU fallback: CPU tensors raise ``Ldmb200Error``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch
import torch.nn as nn

from . import _cabi
from ._cabi import F32, GemmDesc
from .unet import _Act, _Arena, _Plan, _WeightPack, _tc_cpad


# ----------------------------------------------------------------------------------------------
# parameter holders (attribute names = the reference's, model.py / autoencoder.py)
# ----------------------------------------------------------------------------------------------

def _norm(c: int) -> nn.GroupNorm:
    return nn.GroupNorm(32, c, eps=1e-6, affine=True)          # Normalize, model.py:38-39


class ResnetBlock(nn.Module):
    """model.py:82-142 (temb_channels = 0 in the Decoder)."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.norm1 = _norm(cin)
        self.conv1 = nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2 = _norm(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, padding=1)
        if cin != cout:
            self.nin_shortcut = nn.Conv2d(cin, cout, 1)


class AttnBlock(nn.Module):
    """model.py:151-198."""

    def __init__(self, c: int):
        super().__init__()
        self.norm = _norm(c)
        self.q = nn.Conv2d(c, c, 1)
        self.k = nn.Conv2d(c, c, 1)
        self.v = nn.Conv2d(c, c, 1)
        self.proj_out = nn.Conv2d(c, c, 1)


class UpsampleConv(nn.Module):
    """model.py:42-57 (with_conv=True)."""

    _ldmb200_fold_upsample = True   # the weight pack adds the four tap-summed 2x2 weight sets (pack kind 4)

    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)


class Decoder(nn.Module):
    """Parameter layout of model.py:462-533."""

    def __init__(self, *, ch, out_ch, ch_mult=(1, 2, 4, 8), num_res_blocks, attn_resolutions, dropout=0.0,
                 resamp_with_conv=True, in_channels, resolution, z_channels, give_pre_end=False, tanh_out=False,
                 use_linear_attn=False, attn_type="vanilla", **ignorekwargs):
        super().__init__()
        if use_linear_attn or attn_type != "vanilla" or not resamp_with_conv or give_pre_end or tanh_out:
            raise NotImplementedError("only the shipped Decoder variant (vanilla attention, conv upsampling) is built")
        ch_mult = [int(v) for v in ch_mult]
        attn_resolutions = [int(v) for v in attn_resolutions]
        self.num_resolutions = len(ch_mult)
        self.num_res_blocks = int(num_res_blocks)
        block_in = ch * ch_mult[-1]
        curr_res = resolution // 2 ** (self.num_resolutions - 1)
        self.z_channels, self.out_ch = z_channels, out_ch
        self.conv_in = nn.Conv2d(z_channels, block_in, 3, padding=1)
        self.mid = nn.Module()
        self.mid.block_1 = ResnetBlock(block_in, block_in)
        self.mid.attn_1 = AttnBlock(block_in)
        self.mid.block_2 = ResnetBlock(block_in, block_in)
        self.up = nn.ModuleList()
        for i_level in reversed(range(self.num_resolutions)):
            block, attn = nn.ModuleList(), nn.ModuleList()
            block_out = ch * ch_mult[i_level]
            for _ in range(self.num_res_blocks + 1):
                block.append(ResnetBlock(block_in, block_out))
                block_in = block_out
                if curr_res in attn_resolutions:
                    attn.append(AttnBlock(block_in))
            up = nn.Module()
            up.block, up.attn = block, attn
            if i_level != 0:
                up.upsample = UpsampleConv(block_in)
                curr_res *= 2
            self.up.insert(0, up)
        self.norm_out = _norm(block_in)
        self.conv_out = nn.Conv2d(block_in, out_ch, 3, padding=1)


class DownsampleConv(nn.Module):
    """model.py:60-79 (with_conv=True): pad (0,1,0,1) then conv 3x3 stride 2 padding 0."""

    def __init__(self, c: int):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, stride=2, padding=0)


class Encoder(nn.Module):
    """Parameter layout of model.py:368-431."""

    def __init__(self, *, ch, out_ch, ch_mult=(1, 2, 4, 8), num_res_blocks, attn_resolutions, dropout=0.0,
                 resamp_with_conv=True, in_channels, resolution, z_channels, double_z=True, use_linear_attn=False,
                 attn_type="vanilla", **ignore_kwargs):
        super().__init__()
        if use_linear_attn or attn_type != "vanilla" or not resamp_with_conv:
            raise NotImplementedError("only the shipped Encoder variant (vanilla attention, conv downsampling) is built")
        ch_mult = [int(v) for v in ch_mult]
        attn_resolutions = [int(v) for v in attn_resolutions]
        self.num_resolutions = len(ch_mult)
        self.num_res_blocks = int(num_res_blocks)
        self.in_channels = in_channels
        self.out_channels = 2 * z_channels if double_z else z_channels
        self.conv_in = nn.Conv2d(in_channels, ch, 3, padding=1)
        curr_res = resolution
        in_ch_mult = [1] + ch_mult
        self.down = nn.ModuleList()
        block_in = ch
        for i_level in range(self.num_resolutions):
            block, attn = nn.ModuleList(), nn.ModuleList()
            block_in, block_out = ch * in_ch_mult[i_level], ch * ch_mult[i_level]
            for _ in range(self.num_res_blocks):
                block.append(ResnetBlock(block_in, block_out))
                block_in = block_out
                if curr_res in attn_resolutions:
                    attn.append(AttnBlock(block_in))
            down = nn.Module()
            down.block, down.attn = block, attn
            if i_level != self.num_resolutions - 1:
                down.downsample = DownsampleConv(block_in)
                curr_res //= 2
            self.down.append(down)
        self.mid = nn.Module()
        self.mid.block_1 = ResnetBlock(block_in, block_in)
        self.mid.attn_1 = AttnBlock(block_in)
        self.mid.block_2 = ResnetBlock(block_in, block_in)
        self.norm_out = _norm(block_in)
        self.conv_out = nn.Conv2d(block_in, self.out_channels, 3, padding=1)


class _Quantize(nn.Module):
    def __init__(self, n_e: int, e_dim: int):
        super().__init__()
        self.n_e, self.e_dim = n_e, e_dim
        self.embedding = nn.Embedding(n_e, e_dim)


# ----------------------------------------------------------------------------------------------
# the launch plan
# ----------------------------------------------------------------------------------------------

class _VQPlan(_Plan):
    """Static launch list of one ``decode`` for a fixed (B, H, W, precision, quantize)."""

    def __init__(self, model: "VQDecoder", B: int, H: int, W: int, mode: str, quantize: bool):
        self.model, self.B, self.H, self.W, self.L, self.mode = model, B, H, W, 0, mode
        self.quantize = quantize
        self.tc = mode in ("bf16", "f16")
        self.act = {"bf16": _cabi.BF16, "f16": _cabi.F16, "fp32": F32}[mode]
        dev = model._device()
        self.lib = _cabi.load()
        self.ops, self.meta, self.py_ops, self.keep, self.op_outs = [], [], {}, [], {}
        self.arena = _Arena(dev)
        self.pack = model._weight_pack(self.act) if self.tc else None
        f32 = dict(dtype=torch.float32, device=dev)
        up = 2 ** (model.decoder.num_resolutions - 1)
        self.z_in = torch.empty(B, model.embed_dim, H, W, **f32)
        self.indices = torch.empty(B * H * W, dtype=torch.int64, device=dev) if quantize else None
        self.out = torch.empty(B, model.decoder.out_ch, H * up, W * up, **f32)
        self.gn_ws = torch.empty(int(self.lib.ldmb200_groupnorm_ws_floats(B, H * W, 0)), **f32)
        self.graph = None
        self.runs = 0
        self.fold_upsample = os.environ.get("LDMB200_FOLD_UPSAMPLE", "1") == "1"
        self.pinned = set()
        self._build()

    def raw_gemm(self, a_ptr: int, a_dtype: int, k: int, rows: int, w_ptr: int, n: int, bias_ptr: Optional[int],
                 out_ptr: int, out_dtype: int, out_ld: int, what: str):
        """out[rows, n] = A[rows, k] · W[n, k]ᵀ (+ bias[n]) with raw device pointers: both operands may be activations."""
        d = GemmDesc()
        d.a, d.a2, d.c1, d.c2, d.a_dtype = a_ptr, None, k, 0, a_dtype
        d.batch, d.h_in, d.w_in, d.ksize, d.stride, d.upsample = 1, rows, 1, 1, 1, 0
        d.w, d.n, d.bias = w_ptr, n, bias_ptr
        d.rowvec, d.rowvec_ld, d.residual, d.residual_ld, d.geglu = None, 0, None, 0, 0
        d.out, d.out_dtype, d.out_ld, d.out_nchw, d.n_valid = out_ptr, out_dtype, out_ld, 0, n
        self.keep.append(d)
        self.emit("ldmb200_igemm_tc" if self.tc else "ldmb200_igemm_simt", C.byref(d), flops=2.0 * rows * n * k,
                  what=f"M{rows} N{n} K{k} {what}")

    def resnet(self, m: ResnetBlock, name: str, x: _Act, H: int, W: int) -> _Act:
        """model.py:120-142: norm-swish-conv, norm-swish-conv, (nin_shortcut)(x) + h."""
        has_skip = hasattr(m, "nin_shortcut")
        y1, raw = self.groupnorm(x, None, m.norm1, H * W, 1, has_skip)
        h1 = self.gemm(y1, name + ".conv1", m.conv1.weight, m.conv1.bias, H, W, ksize=3, want_stats=True)
        self.free(y1)
        y2, _ = self.groupnorm(h1, None, m.norm2, H * W, 1, False)
        self.free(h1)
        if has_skip:
            res = self.gemm(raw, name + ".nin_shortcut", m.nin_shortcut.weight, m.nin_shortcut.bias, H, W, ksize=1)
            self.free(raw)
        else:
            res = x
        h2 = self.gemm(y2, name + ".conv2", m.conv2.weight, m.conv2.bias, H, W, ksize=3, residual=res, want_stats=True)
        self.free(y2)
        if has_skip:
            self.free(res)
        return h2

    def attn(self, m: AttnBlock, name: str, x: _Act, H: int, W: int) -> _Act:
        """model.py:172-198.  x fp32 [B*N, C] -> x + proj_out(softmax(q kᵀ C^-1/2) v)."""
        B, N, c = self.B, H * W, x.c
        es = 4 if self.act == F32 else 2
        hn, _ = self.groupnorm(x, None, m.norm, N, 0, False)
        q = self.gemm(hn, name + ".q", m.q.weight, m.q.bias, H, W, ksize=1, out_dtype=self.act)
        k = self.gemm(hn, name + ".k", m.k.weight, m.k.bias, H, W, ksize=1, out_dtype=self.act)
        o = self.new(B * N, c, self.act)
        s = self.new(N, N, F32)
        p = self.new(N, N, self.act)
        vt = self.new(c, N, self.act)
        wv = self.pack.ptr(name + ".v") if self.tc else m.v.weight.data_ptr()
        for b in range(B):
            off = b * N * c * es
            self.raw_gemm(q.ptr + off, self.act, c, N, k.ptr + off, N, None, s.ptr, F32, N, "attn q.kT")
            self.emit("ldmb200_softmax_rows", s.ptr, N, p.ptr, self.act, N, N, N, float(int(c) ** (-0.5)),
                      bytes_=N * N * (4 + es), what=f"rows{N} n{N}")
            self.raw_gemm(wv, self.act, c, c, hn.ptr + off, N, None, vt.ptr, self.act, N, "attn vT = Wv.hT")
            self.raw_gemm(p.ptr, self.act, N, N, vt.ptr, c, m.v.bias.data_ptr(), o.ptr + off, self.act, c, "attn P.v")
        self.free(hn, q, k, s, p, vt)
        out = self.gemm(o, name + ".proj_out", m.proj_out.weight, m.proj_out.bias, H, W, ksize=1, residual=x,
                        want_stats=True)
        self.free(o)
        return out

    def up2x(self, m: UpsampleConv, name: str, x: _Act, H: int, W: int) -> _Act:
        """model.py:53-57: nearest 2x, conv 3x3."""
        if self.tc and self.fold_upsample and H & (H - 1) == 0 and W & (W - 1) == 0 and W <= 128:
            # folded into the convolution like the UNet's Upsample (four tap-summed 2x2 convolutions on the low-res map)
            lo = self.new(x.rows, x.c, self.act)
            self.emit("ldmb200_cast_rows", x.ptr, lo.ptr, self.act, x.rows * x.c, bytes_=x.rows * x.c * 6, outs=(lo,))
            out = self.gemm(lo, name + ".conv.fold", m.conv.weight, m.conv.bias, H, W, ksize=3, upsample=1,
                            want_stats=(H * W) % 32 == 0)
            self.free(lo)
            return out
        if self.tc:
            u = self.new(4 * x.rows, x.c, self.act)
            self.emit("ldmb200_upsample2x", x.ptr, u.ptr, self.act, self.B, H, W, x.c,
                      bytes_=x.rows * x.c * (4 + 4 * 2), what=f"rows{x.rows} C{x.c}")
            out = self.gemm(u, name + ".conv", m.conv.weight, m.conv.bias, 2 * H, 2 * W, ksize=3, want_stats=True)
            self.free(u)
            return out
        return self.gemm(x, "", m.conv.weight, m.conv.bias, H, W, ksize=3, upsample=1)

    def _build(self):
        m, B, H, W = self.model, self.B, self.H, self.W
        dec = m.decoder
        zc = dec.z_channels
        cpad = _tc_cpad(zc) if self.tc else zc
        h = self.new(B * H * W, cpad, self.act)
        pq = m.post_quant_conv
        self.emit("ldmb200_vq_lookup_postconv", self.z_in.data_ptr(),
                  m.quantize.embedding.weight.data_ptr() if self.quantize else None, m.quantize.n_e, m.embed_dim,
                  pq.weight.data_ptr(), pq.bias.data_ptr(), zc, h.ptr, self.act, cpad,
                  self.indices.data_ptr() if self.quantize else None, None, B, H * W,
                  flops=(3.0 * m.embed_dim * m.quantize.n_e * B * H * W) if self.quantize else 0.0,
                  what=f"pixels{B * H * W} codes{m.quantize.n_e if self.quantize else 0}")
        x = self.gemm(h, "decoder.conv_in", dec.conv_in.weight, dec.conv_in.bias, H, W, ksize=3, want_stats=True)
        self.free(h)

        def step(fn, *a):
            nonlocal x
            new = fn(*a)
            self.free(x)
            x = new

        step(lambda: self.resnet(dec.mid.block_1, "decoder.mid.block_1", x, H, W))
        step(lambda: self.attn(dec.mid.attn_1, "decoder.mid.attn_1", x, H, W))
        step(lambda: self.resnet(dec.mid.block_2, "decoder.mid.block_2", x, H, W))
        for i_level in reversed(range(dec.num_resolutions)):
            lv = dec.up[i_level]
            for j in range(dec.num_res_blocks + 1):
                step(lambda: self.resnet(lv.block[j], f"decoder.up.{i_level}.block.{j}", x, H, W))
                if len(lv.attn) > 0:
                    step(lambda: self.attn(lv.attn[j], f"decoder.up.{i_level}.attn.{j}", x, H, W))
            if i_level != 0:
                step(lambda: self.up2x(lv.upsample, f"decoder.up.{i_level}.upsample", x, H, W))
                H, W = 2 * H, 2 * W
        y, _ = self.groupnorm(x, None, dec.norm_out, H * W, 1, False)
        self.free(x)
        self.gemm(y, "decoder.conv_out", dec.conv_out.weight, dec.conv_out.bias, H, W, ksize=3,
                  out_nchw_ptr=self.out.data_ptr())
        self.free(y)


class _VQEncPlan(_VQPlan):
    """Static launch list of one ``encode`` (Encoder + quant_conv) for a fixed (B, H, W, precision)."""

    def __init__(self, model: "VQEncoder", B: int, H: int, W: int, mode: str):
        self.model, self.B, self.H, self.W, self.L, self.mode = model, B, H, W, 0, mode
        self.quantize = False
        self.tc = mode in ("bf16", "f16")
        self.act = {"bf16": _cabi.BF16, "f16": _cabi.F16, "fp32": F32}[mode]
        dev = model._device()
        self.lib = _cabi.load()
        self.ops, self.meta, self.py_ops, self.keep, self.op_outs = [], [], {}, [], {}
        self.arena = _Arena(dev)
        self.pack = model._weight_pack(self.act) if self.tc else None
        f32 = dict(dtype=torch.float32, device=dev)
        enc = model.encoder
        f = 2 ** (enc.num_resolutions - 1)
        assert H % f == 0 and W % f == 0, (H, W, f)
        self.x_in = torch.empty(B, enc.in_channels, H, W, **f32)
        self.h_nchw = torch.empty(B, enc.out_channels, H // f, W // f, **f32)      # Encoder output (before quant_conv)
        self.h_nhwc = torch.empty(B * (H // f) * (W // f), model.embed_dim, **f32)
        self.out = torch.empty(B, model.embed_dim, H // f, W // f, **f32)
        self.gn_ws = torch.empty(int(self.lib.ldmb200_groupnorm_ws_floats(B, H * W, 0)), **f32)
        self.graph = None
        self.runs = 0
        self.fold_upsample = os.environ.get("LDMB200_FOLD_UPSAMPLE", "1") == "1"
        self.pinned = set()
        self._build()

    def down2x(self, m: DownsampleConv, name: str, x: _Act, H: int, W: int) -> _Act:
        """model.py:72-76: zero row below / zero column right, conv 3x3 stride 2 (no other padding)."""
        if self.tc:
            sp = self.new(x.rows, x.c, self.act)
            self.emit("ldmb200_space_to_depth2", x.ptr, sp.ptr, self.act, self.B, H, W, x.c,
                      bytes_=x.rows * x.c * (4 + 2), what=f"rows{x.rows} C{x.c}")
            out = self.gemm(sp, name + ".conv", m.conv.weight, m.conv.bias, H, W, ksize=3, stride=2, want_stats=True)
            self.keep[-1].pad_mode = 1
            self.free(sp)
            return out
        out = self.gemm(x, "", m.conv.weight, m.conv.bias, H, W, ksize=3, stride=2)
        self.keep[-1].pad_mode = 1
        return out

    def _build(self):
        m, B, H, W = self.model, self.B, self.H, self.W
        enc = m.encoder
        cin = enc.in_channels
        cpad = _tc_cpad(cin) if self.tc else cin
        h = self.new(B * H * W, cpad, self.act)
        self.emit("ldmb200_nchw_to_nhwc", self.x_in.data_ptr(), h.ptr, B, cin, H * W, cpad, self.act,
                  bytes_=B * H * W * cin * 6, what=f"pixels{B * H * W}")
        x = self.gemm(h, "encoder.conv_in", enc.conv_in.weight, enc.conv_in.bias, H, W, ksize=3, want_stats=True)
        self.free(h)

        def step(fn):
            nonlocal x
            new = fn()
            self.free(x)
            x = new

        for i_level in range(enc.num_resolutions):
            lv = enc.down[i_level]
            for j in range(enc.num_res_blocks):
                step(lambda: self.resnet(lv.block[j], f"encoder.down.{i_level}.block.{j}", x, H, W))
                if len(lv.attn) > 0:
                    step(lambda: self.attn(lv.attn[j], f"encoder.down.{i_level}.attn.{j}", x, H, W))
            if i_level != enc.num_resolutions - 1:
                step(lambda: self.down2x(lv.downsample, f"encoder.down.{i_level}.downsample", x, H, W))
                H, W = H // 2, W // 2
        step(lambda: self.resnet(enc.mid.block_1, "encoder.mid.block_1", x, H, W))
        step(lambda: self.attn(enc.mid.attn_1, "encoder.mid.attn_1", x, H, W))
        step(lambda: self.resnet(enc.mid.block_2, "encoder.mid.block_2", x, H, W))
        y, _ = self.groupnorm(x, None, enc.norm_out, H * W, 1, False)
        self.free(x)
        self.gemm(y, "encoder.conv_out", enc.conv_out.weight, enc.conv_out.bias, H, W, ksize=3,
                  out_nchw_ptr=self.h_nchw.data_ptr())
        self.free(y)
        # quant_conv (autoencoder.py:42,271): 1x1 over z_channels -> embed_dim, fp32, then back to NCHW
        qc = m.quant_conv
        self.emit("ldmb200_vq_lookup_postconv", self.h_nchw.data_ptr(), None, 0, enc.out_channels, qc.weight.data_ptr(),
                  qc.bias.data_ptr(), m.embed_dim, self.h_nhwc.data_ptr(), F32, m.embed_dim, None, None, B, H * W,
                  what=f"quant_conv pixels{B * H * W}")
        self.emit("ldmb200_nhwc_to_nchw", self.h_nhwc.data_ptr(), m.embed_dim, self.out.data_ptr(), B, m.embed_dim, H * W,
                  what=f"pixels{B * H * W}")


# ----------------------------------------------------------------------------------------------
# the drop-in modules
# ----------------------------------------------------------------------------------------------

class VQDecoder(nn.Module):
    """Decode half of ``ldm.models.autoencoder.VQModelInterface`` (ctor: the ``first_stage_config.params`` of the
    shipped YAMLs; ``lossconfig`` / ``ckpt_path`` etc. are accepted and ignored)."""

    def __init__(self, ddconfig, embed_dim, n_embed, lossconfig=None, **ignored):
        super().__init__()
        dd = dict(ddconfig)
        self.embed_dim = int(embed_dim)
        self.decoder = Decoder(**dd)
        self.quantize = _Quantize(int(n_embed), self.embed_dim)
        self.post_quant_conv = nn.Conv2d(self.embed_dim, dd["z_channels"], 1)
        self.precision = os.environ.get("LDMB200_PRECISION", "f16")
        self.use_cuda_graph = os.environ.get("LDMB200_CUDA_GRAPH", "1") != "0"
        self._plans: Dict[tuple, _VQPlan] = {}
        self._packs: Dict[int, _WeightPack] = {}
        for p in self.parameters():
            p.requires_grad_(False)             # the first stage is frozen in every LDM config (ddpm.py:520-525)

    def _device(self):
        return self.post_quant_conv.weight.device

    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        self._plans, self._packs = {}, {}
        return out

    def _weight_pack(self, dtype: int) -> _WeightPack:
        if dtype not in self._packs:
            self._packs[dtype] = _WeightPack(self, dtype)
        return self._packs[dtype]

    def _plan(self, B, H, W, mode, quantize) -> _VQPlan:
        key = (B, H, W, mode, quantize)
        if key not in self._plans:
            if len(self._plans) >= 4:
                self._plans.pop(next(iter(self._plans)))
            self._plans[key] = _VQPlan(self, B, H, W, mode, quantize)
        return self._plans[key]

    @torch.no_grad()
    def decode(self, h: torch.Tensor, force_not_quantize: bool = False) -> torch.Tensor:
        """autoencoder.py:274-283: h [B, embed_dim, H, W] -> image [B, out_ch, H*f, W*f] (fp32, same device)."""
        if not h.is_cuda:
            raise _cabi.Ldmb200Error("dsml_thesis_b200.VQDecoder runs on a CUDA device only (no CPU fallback)")
        mode = self.precision
        if mode not in ("f16", "bf16", "fp32"):
            raise ValueError(f"precision must be 'f16', 'bf16' or 'fp32', got {mode!r}")
        B, E, H, W = h.shape
        assert E == self.embed_dim, (E, self.embed_dim)
        with torch.cuda.device(h.device):
            plan = self._plan(B, H, W, mode, not force_not_quantize)
            plan.z_in.copy_(h)
            stream = torch.cuda.current_stream()
            if plan.pack is not None:
                plan.pack.run(stream.cuda_stream)
            if self.use_cuda_graph and plan.runs >= 1:
                if plan.graph is None:
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr):
                        plan.run(torch.cuda.current_stream().cuda_stream)
                    plan.graph = gr
                plan.graph.replay()
            else:
                plan.run(stream.cuda_stream)
            self.last_indices = plan.indices
            return plan.out.clone()

    def forward(self, h, force_not_quantize: bool = False):
        return self.decode(h, force_not_quantize)


class VQEncoder(nn.Module):
    """Encode half of ``VQModelInterface`` (autoencoder.py:269-272): ``quant_conv(encoder(x))`` — no quantisation, exactly
    like the reference's ``encode``.  Same ctor as ``VQDecoder``; parameters ``encoder.*`` and ``quant_conv.*``."""

    def __init__(self, ddconfig, embed_dim, n_embed=None, lossconfig=None, **ignored):
        super().__init__()
        dd = dict(ddconfig)
        self.embed_dim = int(embed_dim)
        self.encoder = Encoder(**dd)
        self.quant_conv = nn.Conv2d(self.encoder.out_channels, self.embed_dim, 1)
        self.precision = os.environ.get("LDMB200_PRECISION", "f16")
        self.use_cuda_graph = os.environ.get("LDMB200_CUDA_GRAPH", "1") != "0"
        self._plans: Dict[tuple, _VQEncPlan] = {}
        self._packs: Dict[int, _WeightPack] = {}
        for p in self.parameters():
            p.requires_grad_(False)

    def _device(self):
        return self.quant_conv.weight.device

    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        self._plans, self._packs = {}, {}
        return out

    def _weight_pack(self, dtype: int) -> _WeightPack:
        if dtype not in self._packs:
            self._packs[dtype] = _WeightPack(self, dtype)
        return self._packs[dtype]

    @torch.no_grad()
    def encode(self, x: torch.Tensor) -> torch.Tensor:
        """x [B, in_channels, H, W] image in [-1, 1] -> latent [B, embed_dim, H/f, W/f] (fp32, same device)."""
        if not x.is_cuda:
            raise _cabi.Ldmb200Error("dsml_thesis_b200.VQEncoder runs on a CUDA device only (no CPU fallback)")
        mode = self.precision
        if mode not in ("f16", "bf16", "fp32"):
            raise ValueError(f"precision must be 'f16', 'bf16' or 'fp32', got {mode!r}")
        B, Cin, H, W = x.shape
        assert Cin == self.encoder.in_channels, (Cin, self.encoder.in_channels)
        with torch.cuda.device(x.device):
            key = (B, H, W, mode)
            plan = self._plans.get(key)
            if plan is None:
                if len(self._plans) >= 4:
                    self._plans.pop(next(iter(self._plans)))
                plan = self._plans[key] = _VQEncPlan(self, B, H, W, mode)
            plan.x_in.copy_(x)
            stream = torch.cuda.current_stream()
            if plan.pack is not None:
                plan.pack.run(stream.cuda_stream)
            if self.use_cuda_graph and plan.runs >= 1:
                if plan.graph is None:
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr):
                        plan.run(torch.cuda.current_stream().cuda_stream)
                    plan.graph = gr
                plan.graph.replay()
            else:
                plan.run(stream.cuda_stream)
            return plan.out.clone()

    def forward(self, x):
        return self.encode(x)


def _build_encoder_from(first_stage_model) -> VQEncoder:
    enc = first_stage_model.encoder
    nres = enc.num_resolutions
    ch = enc.conv_in.out_channels
    ch_mult = [enc.down[i].block[0].conv1.out_channels // ch for i in range(nres)]
    resolution = getattr(enc, "resolution", 2 ** (nres - 1))
    attn_res = [resolution // 2 ** i for i in range(nres) if len(enc.down[i].attn) > 0]
    zc = enc.conv_out.out_channels
    dd = dict(ch=ch, out_ch=3, ch_mult=ch_mult, num_res_blocks=enc.num_res_blocks, attn_resolutions=attn_res,
              in_channels=enc.conv_in.in_channels, resolution=resolution, z_channels=zc, double_z=False)
    fast = VQEncoder(dd, first_stage_model.quant_conv.out_channels)
    missing, _ = fast.load_state_dict(first_stage_model.state_dict(), strict=False)
    if missing:
        raise _cabi.Ldmb200Error(f"first stage state dict lacks {missing[:4]}...")
    return fast.to(first_stage_model.quant_conv.weight.device)


def accelerated_encode(reference_encode):
    """Class-level wrapper for ``VQModelInterface.encode`` used by ``install()`` (same rules as ``accelerated_decode``)."""
    def encode(self, x):
        if not x.is_cuda or (torch.is_grad_enabled() and x.requires_grad):
            return reference_encode(self, x)
        fast = self.__dict__.get("_ldmb200_encoder")
        fp = _fingerprint(self)
        if fast is None or fast._device() != x.device or self.__dict__.get("_ldmb200_encoder_fp") != fp:
            fast = _build_encoder_from(self)      # first use, device move, or the live weights changed (load_state_dict,
            self.__dict__["_ldmb200_encoder"] = fast   # init_from_ckpt, .half(): new data_ptr / bumped _version)
            self.__dict__["_ldmb200_encoder_fp"] = fp
        return fast.encode(x)
    encode._ldmb200_reference = reference_encode
    return encode


def _fingerprint(module) -> tuple:
    """(storage address, in-place version) of every parameter of the live first stage: changes whenever its weights are
    replaced or overwritten through torch (load_state_dict / init_from_ckpt copy in place and bump ``_version``; ``.half()``
    / ``.to()`` re-allocate), so the private 16-bit copy behind ``install()``'s wrappers is rebuilt instead of going stale."""
    return tuple((p.data_ptr(), p._version) for p in module.parameters())


def _build_from(first_stage_model) -> VQDecoder:
    """A ``VQDecoder`` with the shapes and weights of a live reference ``VQModelInterface`` (read from its own modules)."""
    dec = first_stage_model.decoder
    n_e, e_dim = first_stage_model.quantize.embedding.weight.shape
    nres = dec.num_resolutions
    ch = dec.conv_out.in_channels
    ch_mult = [dec.up[i].block[0].conv1.out_channels // ch for i in range(nres)]
    attn_levels = [i for i in range(nres) if len(dec.up[i].attn) > 0]
    resolution = getattr(dec, "resolution", 2 ** (nres - 1))
    attn_res = [resolution // 2 ** i for i in attn_levels]
    dd = dict(ch=ch, out_ch=dec.conv_out.out_channels, ch_mult=ch_mult, num_res_blocks=dec.num_res_blocks,
              attn_resolutions=attn_res, in_channels=getattr(dec, "in_channels", 3), resolution=resolution,
              z_channels=dec.conv_in.in_channels)
    fast = VQDecoder(dd, e_dim, n_e)
    missing, _ = fast.load_state_dict(first_stage_model.state_dict(), strict=False)
    if missing:
        raise _cabi.Ldmb200Error(f"first stage state dict lacks {missing[:4]}...")
    return fast.to(first_stage_model.post_quant_conv.weight.device)


def accelerate_first_stage(first_stage_model, precision: Optional[str] = None) -> VQDecoder:
    """Builds a ``VQDecoder`` from a live reference ``VQModelInterface`` (weights copied once — the first stage is frozen
    in every LDM config) and rebinds ``first_stage_model.decode`` to it.  Encoding stays on the reference path."""
    fast = _build_from(first_stage_model)
    if precision is not None:
        fast.precision = precision
    first_stage_model.decode = fast.decode
    first_stage_model.__dict__["_ldmb200_decoder"] = fast
    return fast


def accelerated_decode(reference_decode):
    """Class-level wrapper for ``VQModelInterface.decode`` used by ``install()``: CUDA latents under ``no_grad`` go
    through a ``VQDecoder`` built lazily per instance; anything else (CPU tensors, gradients through the decoder) is not
    this path and stays on the reference's own method."""
    def decode(self, h, force_not_quantize=False):
        if not h.is_cuda or (torch.is_grad_enabled() and h.requires_grad):
            return reference_decode(self, h, force_not_quantize)
        fast = self.__dict__.get("_ldmb200_decoder")
        fp = _fingerprint(self)
        if fast is None or fast._device() != h.device or self.__dict__.get("_ldmb200_decoder_fp") != fp:
            fast = _build_from(self)              # first use, device move, or the live weights changed
            self.__dict__["_ldmb200_decoder"] = fast
            self.__dict__["_ldmb200_decoder_fp"] = fp
        return fast.decode(h, force_not_quantize)
    decode._ldmb200_reference = reference_decode
    return decode
