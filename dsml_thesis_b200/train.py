"""DDPM training step on the sm_100a kernels (BASELINE config 4).

Reference: ``LatentDiffusion.p_losses`` ldm/models/diffusion/ddpm.py:1014-1047 (talking_face
ddpm2cond.py:965-998) = ``q_sample`` (ddpm.py:273-276) -> ``apply_model`` -> ``get_loss`` (l2, mean over C,H,W then over
the batch; ``logvar`` is zero and not learned, ``original_elbo_weight`` 0, so ``loss == loss_simple.mean()``), followed by
autograd through ``UNetModel.forward`` (openaimodel.py:710-742), ``configure_optimizers`` AdamW (ddpm.py:1363-1372) and
the ``LitEma`` sweep (ema.py:25-44).  Under DDP the gradients are all-reduced over NCCL.

Here the UNet forward is the same static launch list as sampling, in bf16 operands, except that it keeps what the
backward needs (activations stay in the arena; GEGLU is unfused so its projection is stored; attention also writes the
log-sum-exp).  The backward is a second static launch list, built block by block in reverse:

  * data gradients of every conv / linear run on the SAME tcgen05 implicit-GEMM kernel as the forward, fed with the
    transposed, tap-flipped weight pack (``ldmb200_pack_weights`` kind 3); the stride-2 Downsample conv through a
    zero-inserted gradient map; the residual-stream gradient stays fp32 (fused residual add in the epilogue);
  * weight gradients: ``ldmb200_wgrad`` (bf16 tensor cores, fp32 atomics straight into the nn.Parameter layout);
  * GroupNorm+SiLU / LayerNorm / GEGLU / attention: their own backward kernels (attention recomputes P from the LSE);
  * bias, timestep-embedding and length-1 cross-attention gradients are column sums of the gradient cast.

Parameter gradients land in ONE flat fp32 arena (``p.grad`` are views), so the data-parallel all-reduce is a single
NCCL call per bucket and the fused AdamW+EMA kernel sweeps it in one launch.  There is no PyTorch fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from . import _cabi
from ._cabi import BF16, F32, GemmDesc, LinearBwdJob, PackJob, WgradDesc
from .unet import (BasicTransformerBlock, Downsample, ResBlock, SpatialTransformer, Upsample, _Act, _Buf, _Plan,
                   _tc_cpad, _tc_npad)


# ----------------------------------------------------------------------------------------------
# weights for training: forward operands (GEGLU unpermuted) + the transposed operands of the data gradients
# ----------------------------------------------------------------------------------------------
class _TrainPack:
    def __init__(self, model, dtype: int = BF16):
        dev = model._device()
        self.dtype = dtype
        self.entries: Dict[str, Tuple[int, int, int]] = {}
        jobs: List[tuple] = []
        off = 0

        def add(key, parts, ksize):
            nonlocal off
            c = parts[0].shape[1]
            n_tot = sum(p.shape[0] for p in parts)
            n_pad, c_pad = _tc_npad(n_tot), _tc_cpad(c)
            row = 0
            for i, p in enumerate(parts):
                rows = (n_pad - row) if i == len(parts) - 1 else p.shape[0]
                jobs.append((p, off, 0, p.shape[0], c, ksize, rows, c_pad, row, 32))
                row += p.shape[0]
            self.entries[key] = (off, n_pad, c_pad)
            off += n_pad * ksize * ksize * c_pad * 2
            off = (off + 255) // 256 * 256

        def add_t(key, parts, ksize):
            """kind 3: [rows = pad(c), taps, cols = pad(sum n)] — the B operand of dX = dY W."""
            nonlocal off
            c = parts[0].shape[1]
            n_tot = sum(p.shape[0] for p in parts)
            rows, cols = _tc_npad(c), _tc_npad(n_tot)
            col = 0
            for i, p in enumerate(parts):
                ncols = (cols - col) if i == len(parts) - 1 else p.shape[0]
                jobs.append((p, off, 3, p.shape[0], c, ksize, rows, cols, col, ncols))
                col += p.shape[0]
            self.entries[key + ".T"] = (off, rows, cols)
            off += rows * ksize * ksize * cols * 2
            off = (off + 255) // 256 * 256

        for name, m in model.named_modules():
            if isinstance(m, nn.Conv2d):
                add(name, [m.weight], m.kernel_size[0])
                if name != "input_blocks.0.0":
                    add_t(name, [m.weight], m.kernel_size[0])
            elif isinstance(m, BasicTransformerBlock):
                a = m.attn1
                qkv = [a.to_q.weight, a.to_k.weight, a.to_v.weight]
                add(name + ".attn1.qkv", qkv, 1)
                add_t(name + ".attn1.qkv", qkv, 1)
                for key, lin in ((".attn1.to_out.0", a.to_out[0]), (".ff.net.0.proj", m.ff.net[0].proj),
                                 (".ff.net.2", m.ff.net[2])):
                    add(name + key, [lin.weight], 1)
                    add_t(name + key, [lin.weight], 1)
        self.bf16 = torch.empty(max(off, 256), dtype=torch.uint8, device=dev)
        arr = (PackJob * len(jobs))()
        for j, (p, dst, kind, n, c, ks, n_pad, c_pad, row_off, bn) in zip(arr, jobs):
            j.src, j.dst = p.data_ptr(), self.bf16.data_ptr() + dst
            j.kind, j.n, j.c, j.ksize = kind, n, c, ks
            j.n_pad, j.c_pad, j.row_offset, j.block_n, j.dst_dtype = n_pad, c_pad, row_off, bn, dtype
        self.jobs_dev = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(dev)
        self.n_jobs = len(jobs)

    def ptr(self, key: str) -> int:
        return self.bf16.data_ptr() + self.entries[key][0]

    def run(self, stream: int):
        _cabi.call("ldmb200_pack_weights", self.jobs_dev.data_ptr(), self.n_jobs, stream)


# ----------------------------------------------------------------------------------------------
# per-model training state: the flat gradient arena (p.grad are views of it) and the weight pack
# ----------------------------------------------------------------------------------------------
class _TrainState:
    def __init__(self, model):
        dev = model._device()
        self.params = [p for p in model.parameters()]
        offs, off = {}, 0
        for p in self.params:
            offs[id(p)] = off
            off += (p.numel() + 3) // 4 * 4          # 16-byte aligned slots
        self.numel = off
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.views = {id(p): self.flat[offs[id(p)]: offs[id(p)] + p.numel()].view_as(p) for p in self.params}
        self.offsets = offs
        self.pack = _TrainPack(model)
        self.plans: Dict[tuple, "_TrainPlan"] = {}
        self.anchor = torch.zeros(1, device=dev, requires_grad=True)   # keeps the autograd node alive
        self.launches = 0

    def gptr(self, p) -> int:
        return self.flat.data_ptr() + 4 * self.offsets[id(p)]

    def bucket_bounds(self, model) -> Tuple[int, int]:
        """Element offset where the parameters of output_blocks / out start: [split, numel) is complete once the
        output half of the backward has run (bucket 0 of the gradient all-reduce)."""
        first = next(iter(model.output_blocks.parameters()))
        return self.offsets[id(first)], self.numel


def train_state(model) -> _TrainState:
    if model._train_state is None:
        model._train_state = _TrainState(model)
    return model._train_state


# ----------------------------------------------------------------------------------------------
# the plan: forward that keeps its activations + the backward launch list
# ----------------------------------------------------------------------------------------------
class _TrainPlan(_Plan):
    def __init__(self, model, B: int, H: int, W: int, L: int):
        if L != 1:
            raise NotImplementedError("the training step is built for a length-1 context (every reference config)")
        if model.num_classes is not None:
            raise NotImplementedError("class-conditional (label_emb) training is not built")
        self.state = train_state(model)
        self.gn_info: Dict[int, dict] = {}
        self.records: List[dict] = []
        self.bwd_ops: List[tuple] = []
        self.bwd_meta: List[tuple] = []
        self.building_bwd = False
        self.grad_marks: Dict[int, Tuple[int, int]] = {}   # backward op index -> gradient arena range complete BEFORE it
        self.fwd_graph, self.bwd_graphs, self.steps = None, {}, 0
        self.side_stream = torch.cuda.Stream(device=model._device())
        self.overlap_wgrad = os.environ.get("LDMB200_WGRAD_OVERLAP", "1") == "1"
        super().__init__(model, B, H, W, L, "bf16")
        dev = model._device()
        self.dout_in = torch.empty(B, model.out_channels, H, W, dtype=torch.float32, device=dev)
        self.dctx = torch.zeros(B, 1, model.context_dim, dtype=torch.float32, device=dev)
        self._build_backward()

    def _make_pack(self):
        return self.state.pack

    # forward activations are kept for the backward; gradient buffers are recycled
    def free(self, *acts):
        if self.building_bwd:
            super().free(*acts)

    def emit(self, name: str, *args, flops: float = 0.0, bytes_: float = 0.0, what: str = "", outs=()):
        if not self.building_bwd:
            return super().emit(name, *args, flops=flops, bytes_=bytes_, what=what, outs=outs)
        # weight gradients are leaves of the dependency graph (nothing downstream reads them before the step ends):
        # they run on a side stream and fill the SMs the latency-bound main chain leaves idle
        self.bwd_ops.append((getattr(self.lib, name), args, name, name == "ldmb200_wgrad"))
        self.bwd_meta.append((name, float(flops), float(bytes_), what))

    def g(self, p) -> int:
        return self.state.gptr(p)

    # ---- forward emitters that record what the backward needs ---------------------------------
    def groupnorm(self, x1, x2, gn, hw, silu, want_raw):
        c = x1.c + (x2.c if x2 is not None else 0)
        y = self.new(x1.rows, c, self.act)
        raw = self.new(x1.rows, c, self.act) if want_raw else None
        ws = self.arena.alloc(4 * int(self.lib.ldmb200_groupnorm_ws_floats(self.B, hw, c)))
        s1, s2 = x1.stats, (x2.stats if x2 is not None else None)
        fused = s1 is not None and (x2 is None or s2 is not None)
        if fused:
            self.emit("ldmb200_groupnorm_fused", x1.ptr, x1.c, s1.ptr, x2.ptr if x2 is not None else None,
                      x2.c if x2 is not None else 0, s2.ptr if s2 is not None else None, gn.weight.data_ptr(),
                      gn.bias.data_ptr(), y.ptr, raw.ptr if raw is not None else None, self.act, self.B, hw,
                      gn.num_groups, float(gn.eps), silu, ws.ptr, what=f"rows{x1.rows} C{c} fused")
        else:
            self.emit("ldmb200_groupnorm", x1.ptr, x1.c, x2.ptr if x2 is not None else None,
                      x2.c if x2 is not None else 0, gn.weight.data_ptr(), gn.bias.data_ptr(), y.ptr,
                      raw.ptr if raw is not None else None, self.act, self.B, hw, gn.num_groups, float(gn.eps), silu,
                      ws.ptr, what=f"rows{x1.rows} C{c}")
        self.gn_info[id(y)] = dict(x1=x1, x2=x2, gn=gn, hw=hw, silu=silu, ws=ws, fused=int(fused))
        return y, raw

    def resblock(self, m, name, x1, x2, H, W, emb):
        has_skip = isinstance(m.skip_connection, nn.Conv2d)
        if x2 is not None and not has_skip:
            raise NotImplementedError("identity skip over a concatenated input is not reachable from UNetModel")
        y1, raw = self.groupnorm(x1, x2, m.in_layers[0], H * W, 1, has_skip)
        c0 = m.in_layers[2]
        h1 = self.gemm(y1, name + ".in_layers.2", c0.weight, c0.bias, H, W, ksize=3, rowvec=self.vec[id(m)],
                       want_stats=True)
        y2, _ = self.groupnorm(h1, None, m.out_layers[0], H * W, 1, False)
        if has_skip:
            sc = m.skip_connection
            res = self.gemm(raw, name + ".skip_connection", sc.weight, sc.bias, H * W, 1, ksize=1)
        else:
            res = x1
        c1 = m.out_layers[3]
        h2 = self.gemm(y2, name + ".out_layers.3", c1.weight, c1.bias, H, W, ksize=3, residual=res, want_stats=True)
        self.records.append(dict(kind="res", m=m, name=name, x1=x1, x2=x2, y1=y1, raw=raw, h1=h1, y2=y2, H=H, W=W,
                                 has_skip=has_skip))
        return h2

    def transformer_block(self, m, name, t0, hw, ctx, last):
        Cc, B = t0.c, self.B
        heads, dh = m.attn1.heads, m.attn1.dim_head
        if dh != 32:
            raise NotImplementedError(f"attention kernels are built for d_head = 32 (got {dh})")
        scale = float(m.attn1.scale)
        a = self.layernorm(t0, m.norm1)
        qkv = self.new(t0.rows, 3 * Cc, BF16)
        self.gemm(a, name + ".attn1.qkv", m.attn1.to_q.weight, None, hw, 1, out_dtype=BF16, out=(qkv.ptr, 3 * Cc),
                  n_true=3 * Cc)
        o = self.new(t0.rows, Cc, BF16)
        lse = self.arena.alloc(4 * B * heads * hw)
        self.emit("ldmb200_attention_tc_lse", qkv.ptr, 3 * Cc, 0, Cc, 2 * Cc, o.ptr, Cc, BF16, lse.ptr, B, heads, hw,
                  scale, flops=4.0 * B * heads * hw * hw * dh, what=f"B{B} h{heads} N{hw}")
        lo = m.attn1.to_out[0]
        t2 = self.gemm(o, name + ".attn1.to_out.0", lo.weight, lo.bias, hw, 1, rowvec=self.vec[id(m)], residual=t0)
        f = self.layernorm(t2, m.norm3)
        gp = m.ff.net[0].proj
        inner = gp.weight.shape[0] // 2
        u = self.gemm(f, name + ".ff.net.0.proj", gp.weight, gp.bias, hw, 1, out_dtype=BF16)
        g = self.new(t0.rows, inner, BF16)
        self.emit("ldmb200_geglu_fwd", u.ptr, g.ptr, t0.rows, inner, what=f"rows{t0.rows} inner{inner}")
        l2 = m.ff.net[2]
        t3 = self.gemm(g, name + ".ff.net.2", l2.weight, l2.bias, hw, 1, residual=t2,
                       out_dtype=self.act if last else F32)
        self.records.append(dict(kind="tb", m=m, name=name, t0=t0, a=a, qkv=qkv, o=o, lse=lse, t2=t2, f=f, u=u, g=g,
                                 hw=hw, heads=heads, scale=scale, inner=inner))
        return t3

    def spatial_transformer(self, m, name, x, H, W, ctx):
        y, _ = self.groupnorm(x, None, m.norm, H * W, 0, False)
        t = self.gemm(y, name + ".proj_in", m.proj_in.weight, m.proj_in.bias, H * W, 1)
        nb = len(m.transformer_blocks)
        first = len(self.records)
        for i, blk in enumerate(m.transformer_blocks):
            t = self.transformer_block(blk, f"{name}.transformer_blocks.{i}", t, H * W, ctx, i == nb - 1)
        out = self.gemm(t, name + ".proj_out", m.proj_out.weight, m.proj_out.bias, H * W, 1, residual=x,
                        want_stats=True)
        tbs = self.records[first:]
        del self.records[first:]
        self.records.append(dict(kind="st", m=m, name=name, x=x, y=y, t_last=t, tbs=tbs, hw=H * W))
        return out

    def downsample(self, m, name, x, H, W):
        s = self.new(x.rows, x.c, self.act)
        self.emit("ldmb200_space_to_depth2", x.ptr, s.ptr, self.act, self.B, H, W, x.c)
        out = self.gemm(s, name + ".op", m.op.weight, m.op.bias, H, W, ksize=3, stride=2, want_stats=True)
        self.records.append(dict(kind="down", m=m, name=name, s=s, H=H, W=W, c=x.c))
        return out

    def upsample(self, m, name, x, H, W):
        u = self.new(4 * x.rows, x.c, self.act)
        self.emit("ldmb200_upsample2x", x.ptr, u.ptr, self.act, self.B, H, W, x.c)
        out = self.gemm(u, name + ".conv", m.conv.weight, m.conv.bias, 2 * H, 2 * W, ksize=3, want_stats=True)
        self.records.append(dict(kind="up", m=m, name=name, u=u, H=H, W=W, c=x.c))
        return out

    def run_block(self, block, name, h, x2, H, W, emb, ctx):
        first = len(self.records)
        for j, layer in enumerate(block):
            lname = f"{name}.{j}"
            if isinstance(layer, ResBlock):
                h = self.resblock(layer, lname, h, x2, H, W, emb)
                x2 = None
            elif isinstance(layer, SpatialTransformer):
                h = self.spatial_transformer(layer, lname, h, H, W, ctx)
            elif isinstance(layer, Downsample):
                h = self.downsample(layer, lname, h, H, W)
                H, W = H // 2, W // 2
            elif isinstance(layer, Upsample):
                h = self.upsample(layer, lname, h, H, W)
                H, W = 2 * H, 2 * W
            elif isinstance(layer, nn.Conv2d):
                x_in = h
                h = self.gemm(h, lname, layer.weight, layer.bias, H, W, ksize=3, want_stats=True)
                self.records.append(dict(kind="conv_in", m=layer, name=lname, x=x_in, H=H, W=W))
            else:
                raise NotImplementedError(f"{lname}: {type(layer).__name__}")
        assert x2 is None
        layers = self.records[first:]
        del self.records[first:]
        self.records.append(dict(kind="block", name=name, layers=layers))
        return h, H, W

    def _build(self):
        m, B, H, W = self.model, self.B, self.H, self.W
        self.pinned = set()
        mc, ted = m.model_channels, 4 * m.model_channels
        self.temb = self.new(B, mc, F32)
        self.emit("ldmb200_timestep_embedding", self.t_in.data_ptr(), self.temb.ptr, B, mc, 10000.0)
        self.e1 = self.small_linear(self.temb.ptr, mc, m.time_embed[0], B, 0)
        self.emb = self.small_linear(self.e1.ptr, ted, m.time_embed[2], B, 1)
        ctx = _Act(_Buf(self.ctx_in.view(torch.uint8).view(-1)), B, self.ctx_in.shape[-1], F32)
        self.ctx_act = ctx
        self._prepass_vectors(self.emb, ctx)
        cin = m.in_channels
        h = self.new(B * H * W, _tc_cpad(cin), self.act)
        self.emit("ldmb200_nchw_to_nhwc", self.x_in.data_ptr(), h.ptr, B, cin, H * W, _tc_cpad(cin), self.act)
        hs = []
        for i, blk in enumerate(m.input_blocks):
            h, H, W = self.run_block(blk, f"input_blocks.{i}", h, None, H, W, self.emb, ctx)
            hs.append(h)
        self.n_input = len(m.input_blocks)
        h, H, W = self.run_block(m.middle_block, "middle_block", h, None, H, W, self.emb, ctx)
        for i, blk in enumerate(m.output_blocks):
            h, H, W = self.run_block(blk, f"output_blocks.{i}", h, hs.pop(), H, W, self.emb, ctx)
        self.h_last = h
        self.y_out, _ = self.groupnorm(h, None, m.out[0], H * W, 1, False)
        oc = m.out[2]
        self.gemm(self.y_out, "out.2", oc.weight, oc.bias, H, W, ksize=3, out_nchw_ptr=self.out.data_ptr())

    # ---- backward emitters -------------------------------------------------------------------
    def cast(self, g: _Act, rows_per_sample: int, ps: Optional[Tuple[int, int]] = None, tot: Optional[int] = None,
             tot2: Optional[int] = None, n_valid: int = 0, want16: bool = True) -> Optional[_Act]:
        """fp32 | bf16 gradient -> bf16 GEMM operand, with the per-sample / total column sums on the side."""
        g16 = self.new(g.rows, g.c, BF16) if (want16 and g.dtype == F32) else None
        self.emit("ldmb200_grad_cast", g.ptr, g.dtype, g16.ptr if g16 is not None else None, g.rows, g.c,
                  rows_per_sample, ps[0] if ps else None, ps[1] if ps else 0, tot, tot2, n_valid,
                  bytes_=g.rows * g.c * (4 if g.dtype == F32 else 2), what=f"rows{g.rows} C{g.c}")
        return g16 if g16 is not None else g

    def wgrad(self, dy16: _Act, x16: _Act, weight, h: int, w: int, ksize: int = 1, stride: int = 1, col_off: int = 0,
              n: Optional[int] = None, c: Optional[int] = None):
        d = WgradDesc()
        n = weight.shape[0] if n is None else n
        c = weight.shape[1] if c is None else c
        taps = ksize * ksize
        d.dy, d.x, d.dw = dy16.ptr + 2 * col_off, x16.ptr, self.g(weight)
        d.n_stride, d.c_stride, d.tap_stride = c * taps, taps, 1
        d.dy_ld, d.n, d.c_ld, d.c = dy16.c, n, x16.c, c
        d.batch, d.h, d.w, d.ksize, d.stride = self.B, h, w, ksize, stride
        self.keep.append(d)
        self.pinned.add(id(dy16))     # read asynchronously on the side stream: never recycled by the arena
        self.emit("ldmb200_wgrad", C.byref(d), flops=2.0 * self.B * h * w * n * c * taps,
                  what=f"M{self.B * h * w} N{n} C{c} ks{ksize} s{stride}")

    def dgrad(self, dy16: _Act, key: str, h: int, w: int, ksize: int = 1, out_dtype: int = BF16,
              residual: Optional[_Act] = None, out: Optional[_Act] = None) -> _Act:
        """dX = conv(dY, W^T flipped) on the forward tcgen05 kernel.  Output columns = the layer's input channels."""
        _off, rows, cols = self.pack.entries[key + ".T"]
        assert dy16.c == cols and dy16.dtype == BF16, (key, dy16.c, cols)
        d = GemmDesc()
        d.a, d.a2, d.c1, d.c2, d.a_dtype = dy16.ptr, None, cols, 0, BF16
        d.batch, d.h_in, d.w_in, d.ksize, d.stride, d.upsample = self.B, h, w, ksize, 1, 0
        d.w, d.n = self.pack.ptr(key + ".T"), rows
        d.residual, d.residual_ld = (residual.ptr, residual.c) if residual is not None else (None, 0)
        M = self.B * h * w
        res = out if out is not None else self.new(M, rows, out_dtype)
        assert res.c == rows and res.rows == M and res.dtype == out_dtype
        d.out, d.out_dtype, d.out_ld, d.out_nchw, d.n_valid = res.ptr, out_dtype, rows, 0, rows
        self.keep.append(d)
        self.emit("ldmb200_igemm_tc", C.byref(d), flops=2.0 * M * rows * cols * ksize * ksize,
                  what=f"dgrad M{M} N{rows} K{cols * ksize * ksize} ks{ksize}")
        return res

    def gn_bwd(self, y: _Act, dy16: _Act, addend: Optional[_Act], into: Optional[_Act] = None):
        """-> (dx1, dx2).  ``into``: existing fp32 gradient of x1 to accumulate into (skip tensors of the UNet)."""
        info = self.gn_info[id(y)]
        x1, x2, gn = info["x1"], info["x2"], info["gn"]
        dx1 = into if into is not None else self.new(x1.rows, x1.c, F32)
        dx2 = self.new(x2.rows, x2.c, F32) if x2 is not None else None
        self.emit("ldmb200_groupnorm_bwd", x1.ptr, x1.c, x2.ptr if x2 is not None else None,
                  x2.c if x2 is not None else 0, gn.weight.data_ptr(), gn.bias.data_ptr(), dy16.ptr,
                  addend.ptr if addend is not None else None, addend.dtype if addend is not None else F32, dx1.ptr,
                  1 if into is not None else 0, dx2.ptr if dx2 is not None else None, 0, self.g(gn.weight),
                  self.g(gn.bias), self.B, info["hw"], gn.num_groups, float(gn.eps), info["silu"], info["ws"].ptr,
                  info["fused"], self.gn_ws2.data_ptr(), bytes_=x1.rows * dy16.c * 16, what=f"rows{x1.rows} C{dy16.c}")
        return dx1, dx2

    def ln_bwd(self, x: _Act, ln, dy16: _Act, addend: _Act) -> _Act:
        dx = self.new(x.rows, x.c, F32)
        self.emit("ldmb200_layernorm_bwd", x.ptr, ln.weight.data_ptr(), dy16.ptr, addend.ptr, dx.ptr,
                  self.g(ln.weight), self.g(ln.bias), x.rows, x.c, float(ln.eps), bytes_=x.rows * x.c * 14,
                  what=f"rows{x.rows} C{x.c}")
        return dx

    def queue_small_bwd(self, queue: list, in_ptr, in_ld, weight, bias, dout_ptr, dout_ld, din_ptr, din_ld, rows,
                        silu_in):
        """Defers one small dense layer's backward into a batched launch (``flush_small``); din accumulates."""
        n, k = weight.shape
        queue.append((in_ptr, weight.data_ptr(), dout_ptr, self.g(weight), self.g(bias) if bias is not None else None,
                      din_ptr, in_ld, dout_ld, din_ld, rows, k, n, silu_in))

    def flush_small(self):
        """Two launches: (emb projections + attn2.to_out) then (attn2.to_v, which consumes to_out's input gradient)."""
        for queue, what in ((self.pending_a, "emb_layers + attn2.to_out"), (self.pending_b, "attn2.to_v")):
            if not queue:
                continue
            arr = (LinearBwdJob * len(queue))()
            for j, q in zip(arr, queue):
                (j.in_, j.w, j.dout, j.dw, j.dbias, j.din, j.in_ld, j.dout_ld, j.din_ld, j.rows, j.k, j.n, j.silu_in) = q
            table = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(self.model._device())
            self.keep.append(table)
            self.emit("ldmb200_small_linear_bwd_batched", table.data_ptr(), len(queue), what=f"{what} x{len(queue)}")
            queue.clear()

    def small_bwd(self, in_ptr, in_ld, lin, dout_ptr, dout_ld, din_ptr, din_ld, acc_din, rows, silu_in):
        n, k = lin.weight.shape
        self.emit("ldmb200_small_linear_bwd", in_ptr, in_ld, lin.weight.data_ptr(), dout_ptr, dout_ld,
                  self.g(lin.weight), self.g(lin.bias) if lin.bias is not None else None, din_ptr, din_ld, acc_din, rows,
                  k, n, silu_in, what=f"R{rows} K{k} N{n}")

    # ---- block backward -----------------------------------------------------------------------
    def bwd_resblock(self, r, dout: _Act, into: Optional[_Act]):
        m, name, H, W = r["m"], r["name"], r["H"], r["W"]
        hw = H * W
        c0, c1 = m.in_layers[2], m.out_layers[3]
        sc = m.skip_connection if r["has_skip"] else None
        d2 = self.cast(dout, hw, tot=self.g(c1.bias), tot2=self.g(sc.bias) if sc is not None else None)
        self.wgrad(d2, r["y2"], c1.weight, H, W, 3)
        dy2 = self.dgrad(d2, name + ".out_layers.3", H, W, 3)
        dh1, _ = self.gn_bwd(r["y2"], dy2, None)
        self.free(dy2)
        evp, evld = self.dvec[id(m)]
        d1 = self.cast(dh1, hw, ps=(evp, evld), tot=self.g(c0.bias))
        self.free(dh1)
        # emb_layers = SiLU -> Linear(emb) (openaimodel.py:218-224,264): its output gradient is the per-sample sum
        el = m.emb_layers[1]
        self.queue_small_bwd(self.pending_a, self.emb.ptr, self.emb.c, el.weight, el.bias, evp, evld, self.demb.ptr,
                             self.demb.c, self.B, 1)
        self.wgrad(d1, r["y1"], c0.weight, H, W, 3)
        dy1 = self.dgrad(d1, name + ".in_layers.2", H, W, 3)
        self.free(d1)
        if sc is not None:
            self.wgrad(d2, r["raw"], sc.weight, hw, 1, 1)
            addend = self.dgrad(d2, name + ".skip_connection", hw, 1, 1)
        else:
            addend = dout
        self.free(d2)
        dx1, dx2 = self.gn_bwd(r["y1"], dy1, addend, into)
        self.free(dy1, dout)
        if sc is not None:
            self.free(addend)
        return dx1, dx2

    def bwd_tb(self, r, dt3: _Act) -> _Act:
        m, name, hw = r["m"], r["name"], r["hw"]
        Cc = r["t0"].c
        l2, gp, lo = m.ff.net[2], m.ff.net[0].proj, m.attn1.to_out[0]
        d3 = self.cast(dt3, hw, tot=self.g(l2.bias))
        self.wgrad(d3, r["g"], l2.weight, hw, 1)
        dg = self.dgrad(d3, name + ".ff.net.2", hw, 1)
        self.free(d3)
        du = self.new(dg.rows, 2 * r["inner"], BF16)
        self.emit("ldmb200_geglu_bwd", r["u"].ptr, dg.ptr, du.ptr, dg.rows, r["inner"], what=f"rows{dg.rows}")
        self.free(dg)
        self.cast(du, hw, tot=self.g(gp.bias))
        self.wgrad(du, r["f"], gp.weight, hw, 1)
        df = self.dgrad(du, name + ".ff.net.0.proj", hw, 1)
        self.free(du)
        dt2 = self.ln_bwd(r["t2"], m.norm3, df, dt3)
        self.free(df, dt3)
        # attn2 on a length-1 context is the per-sample row vector to_out(to_v(ctx)): its gradient is the
        # per-sample column sum of dt2; to_q / to_k / norm2 get exactly zero gradient (softmax over one key is 1)
        cvp, cvld = self.dvec[("cv", id(m))]
        d2 = self.cast(dt2, hw, ps=(cvp, cvld), tot=self.g(lo.bias))
        lo2, a2 = m.attn2.to_out[0], m.attn2
        vp, vld = self.vec[("v", id(m))]
        dvp, dvld = self.dvec[("v", id(m))]
        self.queue_small_bwd(self.pending_a, vp, vld, lo2.weight, lo2.bias, cvp, cvld, dvp, dvld, self.B, 0)
        self.queue_small_bwd(self.pending_b, self.ctx_act.ptr, self.ctx_act.c, a2.to_v.weight, None, dvp, dvld,
                             self.dctx.data_ptr(), self.ctx_act.c, self.B, 0)
        self.wgrad(d2, r["o"], lo.weight, hw, 1)
        do = self.dgrad(d2, name + ".attn1.to_out.0", hw, 1)
        self.free(d2)
        dqkv = self.new(do.rows, 3 * Cc, BF16)
        ws = self.arena.alloc(4 * int(self.lib.ldmb200_attention_bwd_ws_floats(self.B, r["heads"], hw)))
        self.emit("ldmb200_attention_bwd", r["qkv"].ptr, 3 * Cc, 0, Cc, 2 * Cc, r["o"].ptr, do.ptr, Cc, r["lse"].ptr,
                  dqkv.ptr, ws.ptr, self.B, r["heads"], hw, r["scale"], flops=10.0 * self.B * r["heads"] * hw * hw * 32,
                  what=f"B{self.B} h{r['heads']} N{hw}")
        self.free(do)
        self.arena.release(ws)
        a1 = m.attn1
        for i, lin in enumerate((a1.to_q, a1.to_k, a1.to_v)):
            self.wgrad(dqkv, r["a"], lin.weight, hw, 1, col_off=i * Cc)
        da = self.dgrad(dqkv, name + ".attn1.qkv", hw, 1)
        self.free(dqkv)
        dt0 = self.ln_bwd(r["t0"], m.norm1, da, dt2)
        self.free(da, dt2)
        return dt0

    def bwd_st(self, r, dout: _Act, into: Optional[_Act]) -> _Act:
        m, name, hw = r["m"], r["name"], r["hw"]
        d16 = self.cast(dout, hw, tot=self.g(m.proj_out.bias))
        self.wgrad(d16, r["t_last"], m.proj_out.weight, hw, 1)
        dt = self.dgrad(d16, name + ".proj_out", hw, 1, out_dtype=F32)
        self.free(d16)
        for tb in reversed(r["tbs"]):
            dt = self.bwd_tb(tb, dt)
        dt16 = self.cast(dt, hw, tot=self.g(m.proj_in.bias))
        self.free(dt)
        self.wgrad(dt16, r["y"], m.proj_in.weight, hw, 1)
        dy = self.dgrad(dt16, name + ".proj_in", hw, 1)
        self.free(dt16)
        dx, _ = self.gn_bwd(r["y"], dy, dout, into)
        self.free(dy, dout)
        return dx

    def bwd_down(self, r, dout: _Act, into: Optional[_Act]) -> _Act:
        m, name, H, W, c = r["m"], r["name"], r["H"], r["W"], r["c"]
        ho, wo = H // 2, W // 2
        d16 = self.cast(dout, ho * wo, tot=self.g(m.op.bias))
        self.free(dout)
        self.wgrad(d16, r["s"], m.op.weight, ho, wo, 3, 2)
        z = self.new(self.B * H * W, c, BF16)
        self.emit("ldmb200_zero_insert2x", d16.ptr, z.ptr, self.B, ho, wo, c)
        self.free(d16)
        dx = self.dgrad(z, name + ".op", H, W, 3, out_dtype=F32, residual=into, out=into)
        self.free(z)
        return dx

    def bwd_up(self, r, dout: _Act) -> _Act:
        m, name, H, W, c = r["m"], r["name"], r["H"], r["W"], r["c"]
        d16 = self.cast(dout, 4 * H * W, tot=self.g(m.conv.bias))
        self.free(dout)
        self.wgrad(d16, r["u"], m.conv.weight, 2 * H, 2 * W, 3)
        du = self.dgrad(d16, name + ".conv", 2 * H, 2 * W, 3, out_dtype=F32)
        self.free(d16)
        dx = self.new(self.B * H * W, c, F32)
        self.emit("ldmb200_downsum2x", du.ptr, dx.ptr, self.B, H, W, c)
        self.free(du)
        return dx

    def bwd_block(self, rec, dout: _Act, into: Optional[_Act]):
        """Backward of one TimestepEmbedSequential; returns (dh, dskip)."""
        layers = rec["layers"]
        d, dskip = dout, None
        for idx in range(len(layers) - 1, -1, -1):
            r = layers[idx]
            tgt = into if idx == 0 else None
            if r["kind"] == "res":
                d, dskip = self.bwd_resblock(r, d, tgt)
            elif r["kind"] == "st":
                d = self.bwd_st(r, d, tgt)
            elif r["kind"] == "down":
                d = self.bwd_down(r, d, tgt)
            elif r["kind"] == "up":
                assert tgt is None
                d = self.bwd_up(r, d)
            elif r["kind"] == "conv_in":
                m = r["m"]
                d16 = self.cast(d, r["H"] * r["W"], tot=self.g(m.bias))
                self.free(d)
                self.wgrad(d16, r["x"], m.weight, r["H"], r["W"], 3)
                self.free(d16)
                d = None
            else:
                raise AssertionError(r["kind"])
        return d, dskip

    def _build_backward(self):
        m, B, H, W = self.model, self.B, self.H, self.W
        self.building_bwd = True
        self.pending_a, self.pending_b = [], []
        dev = m._device()
        self.gn_ws2 = torch.empty(int(self.lib.ldmb200_groupnorm_ws_floats(B, H * W, 0)), dtype=torch.float32, device=dev)
        # per-sample gradient vectors (zeroed at the start of every backward): d emb_out of every ResBlock, d of the
        # cross-attention row vectors and of to_v's outputs, d emb
        res = [mod for _, mod in m.named_modules() if isinstance(mod, ResBlock)]
        tbs = [mod for _, mod in m.named_modules() if isinstance(mod, BasicTransformerBlock)]
        tot_e = sum(mod.out_channels for mod in res)
        tot_v = sum(mod.attn2.to_v.weight.shape[0] for mod in tbs)
        ted = 4 * m.model_channels
        self.scratch = torch.zeros(B * (tot_e + 2 * tot_v + 2 * ted), dtype=torch.float32, device=dev)
        base = self.scratch.data_ptr()
        self.dvec: Dict[object, Tuple[int, int]] = {}
        off = 0
        for mod in res:
            self.dvec[id(mod)] = (base + 4 * off, tot_e)
            off += mod.out_channels
        base_cv, base_v = base + 4 * B * tot_e, base + 4 * B * (tot_e + tot_v)
        off = 0
        for mod in tbs:
            self.dvec[("cv", id(mod))] = (base_cv + 4 * off, tot_v)
            self.dvec[("v", id(mod))] = (base_v + 4 * off, tot_v)
            off += mod.attn2.to_v.weight.shape[0]
        demb_ptr = base + 4 * B * (tot_e + 2 * tot_v)
        self.demb = _Act(_Buf(self.scratch[B * (tot_e + 2 * tot_v): B * (tot_e + 2 * tot_v) + B * ted]), B, ted, F32)
        self.de1 = _Act(_Buf(self.scratch[B * (tot_e + 2 * tot_v) + B * ted:]), B, ted, F32)
        assert self.demb.ptr == demb_ptr
        # ---- out: GN -> SiLU -> conv3x3 (openaimodel.py:682-686)
        oc = m.out[2]
        d16 = self.new(B * H * W, 16, BF16)
        self.emit("ldmb200_nchw_to_nhwc", self.dout_in.data_ptr(), d16.ptr, B, m.out_channels, H * W, 16, BF16)
        self.cast(d16, H * W, tot=self.g(oc.bias), n_valid=m.out_channels)
        self.wgrad(d16, self.y_out, oc.weight, H, W, 3)
        dy = self.dgrad(d16, "out.2", H, W, 3)
        self.free(d16)
        dh, _ = self.gn_bwd(self.y_out, dy, None)
        self.free(dy)
        # ---- output blocks, middle, input blocks in reverse (openaimodel.py:731-737)
        blocks = self.records
        n_in = self.n_input
        dskips: List[_Act] = []
        for rec in reversed(blocks[n_in + 1:]):
            dh, dskip = self.bwd_block(rec, dh, None)
            dskips.append(dskip)          # dskips[j] = gradient of hs[j], the skip popped by output_blocks[n_out-1-j]
        self.flush_small()
        self.grad_marks[len(self.bwd_ops)] = self.state.bucket_bounds(m)
        # middle block: its input is hs[-1], whose gradient buffer already holds the skip contribution
        dh, _ = self.bwd_block(blocks[n_in], dh, dskips[n_in - 1])
        for k in range(n_in - 1, -1, -1):
            # dh is the complete gradient of hs[k]; block k's input is hs[k-1]
            into = dskips[k - 1] if k > 0 else None
            dh, _ = self.bwd_block(blocks[k], dh, into)
        # ---- time_embed: Linear -> SiLU -> Linear (openaimodel.py:506-511)
        self.flush_small()
        te0, te2 = m.time_embed[0], m.time_embed[2]
        self.small_bwd(self.e1.ptr, self.e1.c, te2, self.demb.ptr, ted, self.de1.ptr, ted, 0, B, 1)
        self.small_bwd(self.temb.ptr, self.temb.c, te0, self.de1.ptr, ted, None, 0, 0, B, 0)
        self.building_bwd = False

    # ---- execution -----------------------------------------------------------------------------
    def _run_bwd_range(self, lo: int, hi: int, stream: int):
        cur = torch.cuda.current_stream()
        side = self.side_stream if self.overlap_wgrad else None
        forked = False
        for fn, args, name, on_side in self.bwd_ops[lo:hi]:
            if on_side and side is not None:
                side.wait_stream(cur)            # everything launched so far (the gradient operand) is ordered first
                rc = fn(*args, side.cuda_stream)
                forked = True
            else:
                rc = fn(*args, stream)
            if rc:
                _cabi.check(rc, name)
        if forked:
            cur.wait_stream(side)                # join: parameter gradients complete before the range ends

    def run_forward(self, use_graph: bool):
        """Weight pack + forward launch list; replayed from a CUDA graph after the first eager run (every pointer and
        tensor map of the step is static, so the ~1000 launches cost no host time)."""
        if not use_graph:
            stream = torch.cuda.current_stream().cuda_stream
            self.state.pack.run(stream)
            self.run(stream)
            return
        if self.fwd_graph is None:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                stream = torch.cuda.current_stream().cuda_stream
                self.state.pack.run(stream)
                self.run(stream)
            self.fwd_graph = g
        self.fwd_graph.replay()

    def run_backward(self, stream: int, on_mark=None, use_graph: bool = False):
        """The backward launch list.  With ``on_mark`` (gradient all-reduce buckets) it is cut at the marks so the
        collective of a finished bucket is issued between the two halves."""
        cuts = sorted(self.grad_marks) if on_mark is not None else []
        bounds = [0] + cuts + [len(self.bwd_ops)]
        if not use_graph:
            self.scratch.zero_()
            self.dctx.zero_()
            for k in range(len(bounds) - 1):
                if k > 0:
                    on_mark(*self.grad_marks[bounds[k]])
                self._run_bwd_range(bounds[k], bounds[k + 1], stream)
            return
        key = tuple(bounds)
        graphs = self.bwd_graphs.get(key)
        if graphs is None:
            graphs = []
            for k in range(len(bounds) - 1):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    if k == 0:
                        self.scratch.zero_()
                        self.dctx.zero_()
                    self._run_bwd_range(bounds[k], bounds[k + 1], torch.cuda.current_stream().cuda_stream)
                graphs.append(g)
            self.bwd_graphs[key] = graphs
        for k, g in enumerate(graphs):
            if k > 0:
                on_mark(*self.grad_marks[bounds[k]])
            g.replay()

    def run_timed(self, stream_obj, backward: bool = False):
        if not backward:
            return super().run_timed(stream_obj)
        self.scratch.zero_()
        self.dctx.zero_()
        evs = []
        for fn, args, name, _side in self.bwd_ops:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream_obj)
            rc = fn(*args, stream_obj.cuda_stream)
            e1.record(stream_obj)
            if rc:
                _cabi.check(rc, name)
            evs.append((e0, e1))
        stream_obj.synchronize()
        return [(mt[0], e0.elapsed_time(e1), mt[1], mt[2], mt[3]) for mt, (e0, e1) in zip(self.bwd_meta, evs)]


# ----------------------------------------------------------------------------------------------
# autograd glue
# ----------------------------------------------------------------------------------------------
class _UNetTrainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, context, anchor, model, plan):
        ctx.model, ctx.plan = model, plan
        # the plan owns every saved activation of the step: remember WHICH forward this node belongs to
        ctx.gen, ctx.key = plan.steps, (plan.B, plan.H, plan.W, plan.L)
        return plan.out.clone()

    @staticmethod
    def backward(ctx, dout):
        model, plan = ctx.model, ctx.plan
        st = train_state(model)
        if st.plans.get(ctx.key) is not plan or plan.steps != ctx.gen:
            raise _cabi.Ldmb200Error(
                "dsml_thesis_b200.UNetModel: backward of a training forward whose saved activations were overwritten by "
                "a later training-mode forward (one live forward per model: run forward -> backward -> forward -> "
                "backward; for several loss terms sum them before .backward(), for micro-batches call backward after "
                "each forward)")
        with torch.cuda.device(dout.device):
            stream = torch.cuda.current_stream().cuda_stream
            # per parameter: p.grad is None (zero_grad(set_to_none=True)) -> its slot starts from zero; p.grad already the
            # arena view -> the caller accumulates over micro-batches (every kernel adds); a foreign tensor (someone
            # assigned p.grad) is folded into the slot.  Frozen parameters (requires_grad=False) never get a .grad and do
            # not influence this decision; their slots are scratch.
            live = [p for p in st.params if p.requires_grad]
            is_view = {id(p): (p.grad is not None and p.grad.data_ptr() == st.views[id(p)].data_ptr()) for p in live}
            foreign = [(p, p.grad) for p in live if p.grad is not None and not is_view[id(p)]]
            if not any(is_view.values()):
                st.flat.zero_()                       # the common case: one memset for the whole arena
            else:
                for p in live:
                    if not is_view[id(p)]:
                        st.views[id(p)].zero_()
            plan.dout_in.copy_(dout)
            sync = getattr(model, "_grad_sync", None)
            plan.run_backward(stream, getattr(model, "_grad_bucket_hook", None),
                              use_graph=model.use_cuda_graph and plan.steps > 1)
            st.launches += len(plan.bwd_ops)
            for p, gr in foreign:
                st.views[id(p)].add_(gr.to(st.views[id(p)].dtype))
            if sync is not None and sync.auto:
                sync.finish()                         # gradients are averaged over the ranks before anyone reads .grad
            for p in live:
                p.grad = st.views[id(p)]
            dctx = plan.dctx.clone() if ctx.needs_input_grad[1] else None
        return None, dctx, torch.zeros_like(st.anchor), None, None


def _ensure_grad_sync(model):
    """Multi-rank training without an explicit ``GradAllReduce``: the B200 backward writes ``p.grad`` straight into the flat
    arena, so ``DistributedDataParallel``'s per-parameter autograd hooks never fire for the UNet (the reference's
    ``main.py`` defaults every multi-GPU run to Lightning DDP, main.py:531-543).  Rather than letting the replicas diverge
    silently, the first training forward in an initialised process group attaches an automatic ``GradAllReduce`` (parameters
    broadcast from rank 0 once, like DDP's constructor; gradients averaged at the end of every backward, bucket 0
    overlapped).  ``LDMB200_AUTO_GRAD_SYNC=0`` turns this into an error instead."""
    sync = getattr(model, "_grad_sync", None)
    if sync is not None:
        return sync
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() <= 1:
        return None
    if os.environ.get("LDMB200_AUTO_GRAD_SYNC", "1") == "0":
        raise _cabi.Ldmb200Error("dsml_thesis_b200.UNetModel trains in a %d-rank process group without a GradAllReduce: "
                                 "DDP hooks do not see its gradients; construct dsml_thesis_b200.train.GradAllReduce(unet) "
                                 "or unset LDMB200_AUTO_GRAD_SYNC=0" % dist.get_world_size())
    sync = GradAllReduce(model, auto=True)
    sync.broadcast_parameters(model)
    return sync


def unet_train_forward(model, x, timesteps, context, y=None):
    """``UNetModel.forward`` in training mode: the recording forward + an autograd node for the backward launch list."""
    if not x.is_cuda:
        raise _cabi.Ldmb200Error("dsml_thesis_b200.UNetModel runs on a CUDA device only (no CPU fallback)")
    if model.precision == "fp32":
        raise NotImplementedError("the training step runs on the bf16 tensor-core path only")
    if context is None:
        raise ValueError("context is required (use_spatial_transformer=True, attention.py:170-176)")
    if x.requires_grad:
        raise NotImplementedError("gradients w.r.t. the noisy latent are not built (p_losses does not need them)")
    B, Cin, H, W = x.shape
    assert Cin == model.in_channels and context.dim() == 3 and context.shape[0] == B
    assert timesteps is not None and timesteps.shape == (B,) and not timesteps.dtype.is_floating_point
    st = train_state(model)
    _ensure_grad_sync(model)
    with torch.cuda.device(x.device):
        key = (B, H, W, int(context.shape[1]))
        plan = st.plans.get(key)
        if plan is None:
            st.plans.clear()      # one live training plan: it owns every activation of a step
            plan = _TrainPlan(model, B, H, W, int(context.shape[1]))
            st.plans[key] = plan
        plan.x_in.copy_(x.detach())
        plan.t_in.copy_(timesteps)
        plan.ctx_in.copy_(context.detach())
        plan.steps += 1           # step 1 runs eagerly (lazy per-kernel setup), later steps replay CUDA graphs
        plan.run_forward(model.use_cuda_graph and plan.steps > 1)   # incl. the weight pack: weights change every step
        st.launches += len(plan.ops) + 1
        out = _UNetTrainFn.apply(x, context, st.anchor, model, plan)
    return out if out.dtype == x.dtype else out.to(x.dtype)


class _MSELossFn(torch.autograd.Function):
    """``get_loss(pred, target, mean=False).mean([1,2,3])`` then ``.mean()`` (ddpm.py:1029-1044) in one kernel."""

    @staticmethod
    def forward(ctx, pred, target):
        B = pred.shape[0]
        per = pred[0].numel()
        pred, target = pred.contiguous(), target.contiguous()
        loss_ps = torch.empty(B, dtype=torch.float32, device=pred.device)
        dpred = torch.empty_like(pred)
        _cabi.call("ldmb200_mse_loss_grad", pred.data_ptr(), target.data_ptr(), loss_ps.data_ptr(), dpred.data_ptr(), B,
                   per, torch.cuda.current_stream().cuda_stream)
        ctx.save_for_backward(dpred)
        ctx.mark_non_differentiable(loss_ps)
        return loss_ps.mean(), loss_ps

    @staticmethod
    def backward(ctx, gl, _gps):
        (dpred,) = ctx.saved_tensors
        return dpred * gl, None


def q_sample(handle, x_start, t, noise):
    """ddpm.py:273-276 as one fused kernel (bit-identical to the reference's extract_into_tensor arithmetic)."""
    x_start, noise = x_start.contiguous(), noise.contiguous()
    out = torch.empty_like(x_start)
    _cabi.call("ldmb200_q_sample", x_start.data_ptr(), noise.data_ptr(), t.data_ptr(),
               handle.sqrt_alphas_cumprod.data_ptr(), handle.sqrt_one_minus_alphas_cumprod.data_ptr(), out.data_ptr(),
               x_start.shape[0], x_start[0].numel(), torch.cuda.current_stream().cuda_stream)
    return out


def p_losses(handle, x_start, cond, t, noise=None, c_concat=None):
    """``LatentDiffusion.p_losses`` (ddpm.py:1014-1047; TF ddpm2cond.py:965-998) for parameterization "eps", l2 loss,
    zero fixed logvar, l_simple_weight 1, original_elbo_weight 0.  Returns ``(loss, loss_dict)`` like the reference."""
    noise = torch.randn_like(x_start) if noise is None else noise
    x_noisy = q_sample(handle, x_start, t, noise)
    model_output = handle.apply_model(x_noisy, t, cond, c_concat) if handle.two_cond else handle.apply_model(x_noisy, t, cond)
    loss, loss_ps = _MSELossFn.apply(model_output.float(), noise)
    prefix = "train" if handle.training else "val"
    loss_dict = {f"{prefix}/loss_simple": loss_ps.mean().detach(), f"{prefix}/loss_gamma": loss.detach(),
                 f"{prefix}/loss": loss.detach()}
    return loss, loss_dict


# ----------------------------------------------------------------------------------------------
# optimizer side: fused AdamW + EMA over the flat arenas; data-parallel gradient all-reduce
# ----------------------------------------------------------------------------------------------
class FusedAdamWEMA:
    """AdamW (ddpm.py:1363-1372: lr, default betas/eps/weight_decay 1e-2) + ``LitEma`` (ema.py:25-44, decay 0.9999 with
    the ``(1+n)/(10+n)`` warm-up) in ONE launch over the UNet's flat parameter / gradient arenas.  The parameters are
    re-pointed into a flat fp32 arena (same values, same ``nn.Parameter`` objects)."""

    def __init__(self, model, lr: float, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2,
                 ema_decay: Optional[float] = 0.9999):
        st = train_state(model)
        self.model, self.st = model, st
        self.lr, self.betas, self.eps, self.wd = lr, betas, eps, weight_decay
        self.pflat = torch.zeros_like(st.flat)
        for p in st.params:
            v = self.pflat[st.offsets[id(p)]: st.offsets[id(p)] + p.numel()].view_as(p)
            v.copy_(p.data)
            p.data = v
        model._plans = {}            # raw parameter pointers changed
        model._packs = {}
        model._pack_valid = set()
        st.plans.clear()
        st.pack = _TrainPack(model)
        self.m = torch.zeros_like(st.flat)
        self.v = torch.zeros_like(st.flat)
        self.ema = self.pflat.clone() if ema_decay is not None else None
        self.ema_decay, self.step_n = ema_decay, 0

    def step(self):
        self.step_n += 1
        b1, b2 = self.betas
        omd = 0.0
        if self.ema is not None:
            decay = min(self.ema_decay, (1 + self.step_n) / (10 + self.step_n))
            omd = 1.0 - decay
        _cabi.call("ldmb200_adamw_ema", self.pflat.data_ptr(), self.st.flat.data_ptr(), self.m.data_ptr(),
                   self.v.data_ptr(), self.ema.data_ptr() if self.ema is not None else None, self.st.numel, self.lr, b1,
                   b2, self.eps, self.wd, 1 - b1 ** self.step_n, 1 - b2 ** self.step_n, omd,
                   torch.cuda.current_stream().cuda_stream)

    def zero_grad(self):
        for p in self.st.params:
            p.grad = None


class GradAllReduce:
    """Data-parallel gradient averaging over NCCL / NVLink (the one collective of the training path): the
    flat arena is reduced in two buckets — the output half of the UNet as soon as its backward is done (overlapping
    the rest of the backward on a side stream), the rest at the end.  Constructing it hooks the bucket into the
    model's backward; ``finish()`` must be called before the optimizer step.  ``flat`` may be given directly (CPU /
    gloo tests of the bucket logic)."""

    def __init__(self, model=None, group=None, flat: Optional[torch.Tensor] = None,
                 wire_dtype: Optional[torch.dtype] = None, auto: bool = False):
        """``wire_dtype=torch.bfloat16`` halves the NVLink bytes (SURVEY §8e "bf16 grads optional"): each bucket is cast to a
        bf16 staging buffer, pre-divided by the world size (so the sum cannot overflow where fp32 would not), summed in
        bf16 by the collective and cast back into the fp32 arena.  Default: fp32 on the wire, exact."""
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.world = dist.get_world_size(group)
        self.flat = flat if flat is not None else train_state(model).flat
        self.side = torch.cuda.Stream(device=self.flat.device) if self.flat.is_cuda else None
        self.wire_dtype = wire_dtype if wire_dtype not in (None, torch.float32) else None
        self.wire = torch.empty_like(self.flat, dtype=self.wire_dtype) if self.wire_dtype is not None else None
        self.split = None
        self.auto = auto      # True: the backward node itself calls finish() (no explicit sync.finish() in user code)
        if model is not None:
            model._grad_bucket_hook = self.bucket_ready
            model._grad_sync = self

    def broadcast_parameters(self, model, src: int = 0):
        """Rank ``src``'s parameters to every rank, once (what ``DistributedDataParallel.__init__`` does for the parameters
        it manages; the UNet's are excluded from DDP by ``install()``)."""
        with torch.no_grad():
            for p in model.parameters():
                self.dist.broadcast(p.data, src=src, group=self.group)
        if hasattr(model, "invalidate_weight_cache"):
            model.invalidate_weight_cache()

    def _reduce(self, lo: int, hi: int):
        if self.wire is None:
            self.dist.all_reduce(self.flat[lo:hi], group=self.group)
            return
        w = self.wire[lo:hi]
        torch.mul(self.flat[lo:hi], 1.0 / self.world, out=w)      # fp32 product, rounded once to the wire dtype
        self.dist.all_reduce(w, group=self.group)
        self.flat[lo:hi].copy_(w)

    def bucket_ready(self, lo: int, hi: int):
        """Elements [lo, hi) of the arena are final: start their all-reduce while the backward continues."""
        self.split = lo
        if self.side is None:
            self._reduce(lo, hi)
            return
        self.side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.side):
            self._reduce(lo, hi)

    def finish(self):
        lo = self.split if self.split is not None else self.flat.numel()
        if lo > 0:
            self._reduce(0, lo)
        if self.side is not None:
            torch.cuda.current_stream().wait_stream(self.side)
        if self.wire is None:
            self.flat.mul_(1.0 / self.world)
        self.split = None
