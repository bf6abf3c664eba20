"""``UNetModel`` — drop-in for ``ldm.modules.diffusionmodules.openaimodel.UNetModel``
(reference openaimodel.py:413-742) whose forward runs entirely on the sm_100a kernels of
``libldmb200.so``.

What is kept from the reference interface (SURVEY §8b "UNet contract"):
  * constructor keyword arguments and their meaning (openaimodel.py:443-469), incl. OmegaConf
    ``ListConfig`` values;
  * ``forward(x, timesteps=None, context=None, y=None, **kwargs)`` -> tensor of ``x.dtype`` and
    shape ``[N, out_channels, H, W]`` (openaimodel.py:710-742);
  * the parameter inventory: same 626 names / shapes / fp32 ``nn.Parameter`` s and no buffers, so
    checkpoints (``load_state_dict(strict=False)``), ``LitEma`` and AdamW work unchanged;
  * initialisation, incl. the 34 ``zero_module`` layers (openaimodel.py:229-231,685;
    attention.py:244-248).

What is different: the modules below only HOLD parameters.  ``forward`` compiles, per
(batch, H, W, context length, precision), a static list of C-ABI kernel launches over a channels-last
activation arena (``_Plan``) and replays it; optionally through a CUDA graph.

Precision modes (``LDMB200_PRECISION`` or ``model.precision``):
  * ``"f16"`` (default) — tcgen05 implicit GEMM + flash attention with 16-bit operands, fp32 accumulation in
    tensor memory, fp32 residual stream, fp32 norm statistics and softmax.  GEMM operands (weights, GroupNorm /
    LayerNorm / GEGLU / attention outputs) are IEEE fp16 with saturating conversions: same tensor-core rate as
    bf16 but 3 more mantissa bits, which is what brings the per-step eps error from ~1.2e-2 (pure bf16 — also
    what the reference's own bf16 autocast gives, SURVEY hard part 2) down to ~1e-3.  Attention internals
    (q, k, v, P) stay bf16 so that scores keep the fp32 exponent range;
  * ``"bf16"`` — the same kernels with every 16-bit operand in bf16;
  * ``"fp32"`` — fp32 SIMT kernels (parity mode, per-step eps rel-L2 <= 1e-4).
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from . import _cabi
from ._cabi import BF16, F16, F32, GemmDesc, LinearJob, PackJob


# ----------------------------------------------------------------------------------------------
# parameter-holder modules: same attribute names as the reference so state_dict keys match
# ----------------------------------------------------------------------------------------------

def _zero(m: nn.Module) -> nn.Module:
    for p in m.parameters():
        p.detach().zero_()
    return m


class _Slot(nn.Identity):
    """Place-holder keeping the reference's Sequential indices (SiLU / Dropout have no parameters)."""


class ResBlock(nn.Module):
    """Parameters of openaimodel.py:163-241 (no up/down, no scale-shift norm)."""

    def __init__(self, channels: int, emb_channels: int, out_channels: int):
        super().__init__()
        self.channels, self.out_channels = channels, out_channels
        self.in_layers = nn.Sequential(nn.GroupNorm(32, channels), _Slot(),
                                       nn.Conv2d(channels, out_channels, 3, padding=1))
        self.emb_layers = nn.Sequential(_Slot(), nn.Linear(emb_channels, out_channels))
        self.out_layers = nn.Sequential(nn.GroupNorm(32, out_channels), _Slot(), _Slot(),
                                        _zero(nn.Conv2d(out_channels, out_channels, 3, padding=1)))
        if out_channels == channels:
            self.skip_connection = nn.Identity()
        else:
            self.skip_connection = nn.Conv2d(channels, out_channels, 1)


class CrossAttention(nn.Module):
    """Parameters of attention.py:152-168."""

    def __init__(self, query_dim: int, context_dim: Optional[int], heads: int, dim_head: int):
        super().__init__()
        inner = heads * dim_head
        context_dim = query_dim if context_dim is None else context_dim
        self.heads, self.dim_head, self.scale = heads, dim_head, dim_head ** -0.5
        self.to_q = nn.Linear(query_dim, inner, bias=False)
        self.to_k = nn.Linear(context_dim, inner, bias=False)
        self.to_v = nn.Linear(context_dim, inner, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, query_dim), _Slot())


class GEGLU(nn.Module):
    def __init__(self, dim_in: int, dim_out: int):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)


class FeedForward(nn.Module):
    """Parameters of attention.py:47-64 with ``glu=True``, ``mult=4``."""

    def __init__(self, dim: int, mult: int = 4):
        super().__init__()
        self.net = nn.Sequential(GEGLU(dim, dim * mult), _Slot(), nn.Linear(dim * mult, dim))


class BasicTransformerBlock(nn.Module):
    """Parameters of attention.py:196-207."""

    def __init__(self, dim: int, n_heads: int, d_head: int, context_dim: Optional[int]):
        super().__init__()
        self.attn1 = CrossAttention(dim, None, n_heads, d_head)
        self.ff = FeedForward(dim)
        self.attn2 = CrossAttention(dim, context_dim, n_heads, d_head)
        self.norm1 = nn.LayerNorm(dim)
        self.norm2 = nn.LayerNorm(dim)
        self.norm3 = nn.LayerNorm(dim)


class SpatialTransformer(nn.Module):
    """Parameters of attention.py:218-248."""

    def __init__(self, in_channels: int, n_heads: int, d_head: int, depth: int, context_dim: Optional[int]):
        super().__init__()
        inner = n_heads * d_head
        self.in_channels, self.n_heads, self.d_head = in_channels, n_heads, d_head
        self.norm = nn.GroupNorm(32, in_channels, eps=1e-6)
        self.proj_in = nn.Conv2d(in_channels, inner, 1)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner, n_heads, d_head, context_dim) for _ in range(depth)])
        self.proj_out = _zero(nn.Conv2d(inner, in_channels, 1))


class Downsample(nn.Module):
    """openaimodel.py:134-160 with ``use_conv=True``."""

    def __init__(self, channels: int):
        super().__init__()
        self.channels = channels
        self.op = nn.Conv2d(channels, channels, 3, stride=2, padding=1)


class Upsample(nn.Module):
    """openaimodel.py:91-119 with ``use_conv=True``."""

    def __init__(self, channels: int):
        super().__init__()
        self.channels = channels
        self.conv = nn.Conv2d(channels, channels, 3, padding=1)


class TimestepEmbedSequential(nn.Sequential):
    """Container with the reference's name (openaimodel.py:74-88); children are parameter holders."""


def _aslist(v):
    """OmegaConf ListConfig / tuple / set -> list (openaimodel.py:474-478)."""
    if v is None or isinstance(v, (int, float)):
        return v
    return list(v)


# ----------------------------------------------------------------------------------------------
# tensor-core weight pack: every conv / linear weight as bf16 [n_pad, taps, c_pad] (K-major)
# ----------------------------------------------------------------------------------------------

def _tc_npad(n: int) -> int:
    return 16 if n <= 16 else (n + 31) // 32 * 32


def _tc_cpad(c: int) -> int:
    return (c + 31) // 32 * 32


def _tc_bn(n: int, geglu: bool = False) -> int:
    """N tile the C side picks for a packed weight with ``n`` rows (igemm_tc.cu): GEGLU projections run on
    128-wide tiles with 16 epilogue warps (one 32-column chunk per warp), everything else on 160 / 32 / 16."""
    if geglu and n % 128 == 0:
        return 128
    return 160 if n % 160 == 0 else (32 if n % 32 == 0 else 16)


class _WeightPack:
    """bf16 K-major copies of the UNet's weights for ``ldmb200_igemm_tc`` plus the device job table that
    (re)fills them in ONE launch (``ldmb200_pack_weights``).  Layout depends on the model only."""

    def __init__(self, model: "UNetModel", dtype: int = BF16):
        dev = model._device()
        self.dtype = dtype
        self.entries: Dict[str, Tuple[int, int, int]] = {}   # key -> (byte offset, n_pad, c_pad)
        self.f32_entries: Dict[str, int] = {}                 # key -> float offset (GEGLU biases)
        self.nbytes: Dict[str, int] = {}                      # key -> bytes of the packed matrix
        jobs: List[tuple] = []
        off = 0
        foff = 0

        def add(key, parts, ksize, kind=0):
            """parts: list of weight Parameters stacked along n (fused QKV)."""
            nonlocal off
            c = parts[0].shape[1]
            n_tot = sum(p.shape[0] for p in parts)
            n_pad, c_pad = _tc_npad(n_tot), _tc_cpad(c)
            base = off
            row = 0
            for i, p in enumerate(parts):
                last = i == len(parts) - 1
                rows = (n_pad - row) if last else p.shape[0]
                jobs.append((p, base, kind, p.shape[0], c, ksize, rows, c_pad, row, _tc_bn(n_pad, kind == 1)))
                row += p.shape[0]
            self.entries[key] = (base, n_pad, c_pad)
            self.nbytes[key] = n_pad * ksize * ksize * c_pad * 2
            off += n_pad * ksize * ksize * c_pad * 2
            off = (off + 255) // 256 * 256

        def add_fold(key, w):
            """Upsample.conv folded with the nearest-2x upsample (K10): four tap-summed 2x2 weight sets (pack kind 4)."""
            nonlocal off
            n, c = w.shape[0], w.shape[1]
            n_pad, c_pad = _tc_npad(n), _tc_cpad(c)
            jobs.append((w, off, 4, n, c, 3, n_pad, c_pad, 0, _tc_bn(n_pad)))
            self.entries[key] = (off, n_pad, c_pad)
            self.nbytes[key] = 4 * n_pad * 4 * c_pad * 2
            off += 4 * n_pad * 4 * c_pad * 2
            off = (off + 255) // 256 * 256

        for name, m in model.named_modules():
            if isinstance(m, Upsample) or getattr(m, "_ldmb200_fold_upsample", False):   # (first-stage UpsampleConv too)
                add_fold(name + ".conv.fold", m.conv.weight)
            if isinstance(m, nn.Conv2d):
                add(name, [m.weight], m.kernel_size[0])
            elif isinstance(m, BasicTransformerBlock):
                a = m.attn1
                add(name + ".attn1.qkv", [a.to_q.weight, a.to_k.weight, a.to_v.weight], 1)
                add(name + ".attn1.to_out.0", [a.to_out[0].weight], 1)
                add(name + ".attn2.to_q", [m.attn2.to_q.weight], 1)
                g = m.ff.net[0].proj
                add(name + ".ff.net.0.proj", [g.weight], 1, kind=1)
                self.f32_entries[name + ".ff.net.0.proj.bias"] = foff
                jobs.append((g.bias, ("f32", foff), 2, g.bias.shape[0], 1, 1, g.bias.shape[0], 1, 0,
                             _tc_bn(g.weight.shape[0], True)))
                foff += g.bias.shape[0]
                add(name + ".ff.net.2", [m.ff.net[2].weight], 1)
        self.bf16 = torch.empty(max(off, 256), dtype=torch.uint8, device=dev)
        self.f32 = torch.empty(max(foff, 1), dtype=torch.float32, device=dev)
        arr = (PackJob * len(jobs))()
        for i, (p, dst, kind, n, c, ks, n_pad, c_pad, row_off, bn) in enumerate(jobs):
            j = arr[i]
            j.src = p.data_ptr()
            if isinstance(dst, tuple):
                j.dst = self.f32.data_ptr() + 4 * dst[1]
            else:
                j.dst = self.bf16.data_ptr() + dst
            j.kind, j.n, j.c, j.ksize = kind, n, c, ks
            j.n_pad, j.c_pad, j.row_offset, j.block_n = n_pad, c_pad, row_off, bn
            j.dst_dtype = dtype
        host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
        self.jobs_dev = host.to(dev)
        self.n_jobs = len(jobs)
        self.param_ptrs = [p.data_ptr() for (p, *_r) in jobs]

    def ptr(self, key: str) -> int:
        return self.bf16.data_ptr() + self.entries[key][0]

    def fptr(self, key: str) -> int:
        return self.f32.data_ptr() + 4 * self.f32_entries[key]

    def run(self, stream: int):
        _cabi.call("ldmb200_pack_weights", self.jobs_dev.data_ptr(), self.n_jobs, stream)


# ----------------------------------------------------------------------------------------------
# activation arena: plan-time allocator with reuse (all launches are ordered on one stream)
# ----------------------------------------------------------------------------------------------

class _Buf:
    __slots__ = ("t", "ptr", "nbytes")

    def __init__(self, t: torch.Tensor):
        self.t, self.ptr, self.nbytes = t, t.data_ptr(), t.numel() * t.element_size()


class _Arena:
    def __init__(self, device):
        self.device = device
        self.free: List[_Buf] = []
        self.all: List[_Buf] = []

    def alloc(self, nbytes: int) -> _Buf:
        nbytes = max(256, (int(nbytes) + 255) // 256 * 256)
        best = None
        for b in self.free:
            if b.nbytes >= nbytes and (best is None or b.nbytes < best.nbytes):
                best = b
        if best is not None and best.nbytes <= 2 * nbytes:
            self.free.remove(best)
            return best
        b = _Buf(torch.empty(nbytes, dtype=torch.uint8, device=self.device))
        self.all.append(b)
        return b

    def release(self, b: Optional[_Buf]):
        if b is not None:
            self.free.append(b)

    def total_bytes(self) -> int:
        return sum(b.nbytes for b in self.all)


class _Act:
    """A channels-last activation [rows, c] living in an arena buffer."""
    __slots__ = ("buf", "rows", "c", "dtype", "stats")

    def __init__(self, buf: _Buf, rows: int, c: int, dtype: int):
        self.buf, self.rows, self.c, self.dtype = buf, rows, c, dtype
        self.stats: Optional[_Buf] = None   # GroupNorm column statistics left by the producing GEMM epilogue

    @property
    def ptr(self) -> int:
        return self.buf.ptr


# ----------------------------------------------------------------------------------------------
# execution plan: the UNet forward (openaimodel.py:710-742) as a static list of C-ABI launches
# ----------------------------------------------------------------------------------------------

class _Plan:
    def __init__(self, model: "UNetModel", B: int, H: int, W: int, L: int, mode: str):
        self.model, self.B, self.H, self.W, self.L, self.mode = model, B, H, W, L, mode
        self.tc = mode in ("bf16", "f16")
        self.act = {"bf16": BF16, "f16": F16, "fp32": F32}[mode]   # dtype of GEMM A operands
        self.qkv_dt = BF16 if self.tc else F32                     # attention_tc takes bf16 q/k/v
        dev = model._device()
        self.lib = _cabi.load()
        self.ops: List[tuple] = []       # (c function, args without the trailing stream)
        self.meta: List[tuple] = []      # (kernel name, algorithmic flops, algorithmic bytes) per op
        self.py_ops: Dict[int, object] = {}  # op index -> python callable run before that op
        self.op_outs: Dict[int, list] = {}   # op index -> [(buffer, elements, dtype)] 16-bit outputs of that op
        self.keep: List[object] = []     # ctypes structs referenced by pointer from ops
        self.arena = _Arena(dev)
        self.pack = self._make_pack()
        f32 = dict(dtype=torch.float32, device=dev)
        self.x_in = torch.empty(B, model.in_channels, H, W, **f32)
        self.t_in = torch.empty(B, dtype=torch.int64, device=dev)
        self.ctx_in = torch.empty(B, max(L, 1), model.context_dim or 1, **f32)
        self.y_in = torch.empty(B, dtype=torch.int64, device=dev) if model.num_classes is not None else None
        self.out = torch.empty(B, model.out_channels, H, W, **f32)
        self.gn_ws = torch.empty(int(self.lib.ldmb200_groupnorm_ws_floats(B, H * W, 0)), **f32)
        # split-K scratch of the tensor-core GEMMs (ldmb200_gemm_desc.splitk_ws): arrival counters (zero) + fp32 partials.
        # The launcher only splits grids of <= 74 tiles with K >= 2560 (batch 1-4 sampling); UNetModel.split_k turns it on
        self.splitk_ws = None
        if self.tc and getattr(model, "split_k", False):
            self.splitk_ws = torch.zeros((_cabi.SPLITK_COUNTER_BYTES + (16 << 20)) // 4, **f32)
        self.graph = None
        self.runs = 0
        self.fold_upsample = getattr(self, "fold_upsample", os.environ.get("LDMB200_FOLD_UPSAMPLE", "1") == "1")
        self.tc_gemms: List[tuple] = []   # (desc, bytes of its packed weights, GEMM rows) in launch order
        self._build()
        self._link_weight_prefetch()

    def _link_weight_prefetch(self):
        """Small-batch evals read every GEMM's weights once, cold, through a few CTAs: each tensor-core GEMM pulls the
        weights of the NEXT one into L2 while it runs (``ldmb200_gemm_desc.prefetch``) when that next GEMM has at most
        ``LDMB200_WPREFETCH_ROWS`` rows — above that its many M tiles share the weights through L2 anyway."""
        max_rows = int(os.environ.get("LDMB200_WPREFETCH_ROWS", "16384"))
        for (d, _, _), (nxt, nbytes, rows) in zip(self.tc_gemms, self.tc_gemms[1:]):
            if nbytes and rows <= max_rows and nbytes % 16 == 0:
                d.prefetch, d.prefetch_bytes = nxt.w, nbytes

    # ---- small helpers -------------------------------------------------------------------
    def _make_pack(self):
        return self.model._weight_pack(self.act) if self.tc else None

    def new(self, rows: int, c: int, dtype: int) -> _Act:
        return _Act(self.arena.alloc(rows * c * (4 if dtype == F32 else 2)), rows, c, dtype)

    def free(self, *acts):
        for a in acts:
            if a is not None and id(a) not in self.pinned:
                self.arena.release(a.buf)
                if a.stats is not None:
                    self.arena.release(a.stats)
                    a.stats = None

    def emit(self, name: str, *args, flops: float = 0.0, bytes_: float = 0.0, what: str = "", outs=()):
        """Appends one launch.  ``flops`` / ``bytes_`` are the ALGORITHMIC work of the launch (no padding, no
        re-reads) used by ``bench.py`` for the roofline; they do not influence execution.  ``outs``: the 16-bit
        activations the launch writes (``f16_saturation_report`` inspects them right after the launch)."""
        self.op_outs[len(self.ops)] = [(o.buf, o.rows * o.c, o.dtype) for o in outs if o is not None and o.dtype != F32]
        self.ops.append((getattr(self.lib, name), args, name))
        self.meta.append((name, float(flops), float(bytes_), what))

    def view_f32(self, a: _Act) -> torch.Tensor:
        return a.buf.t[: a.rows * a.c * 4].view(torch.float32).view(a.rows, a.c)

    # ---- kernel emitters -----------------------------------------------------------------
    def small_linear(self, in_ptr: int, in_ld: int, lin: nn.Linear, rows: int, silu_in: int) -> _Act:
        n, k = lin.weight.shape
        o = self.new(rows, n, F32)
        self.emit("ldmb200_small_linear", in_ptr, in_ld, lin.weight.data_ptr(),
                  lin.bias.data_ptr() if lin.bias is not None else None, o.ptr, n, rows, k, n, silu_in,
                  what=f"R{rows} K{k} N{n}")
        return o

    def small_linear_batched(self, jobs: List[tuple], what: str):
        """One launch for a table of independent small dense layers.  jobs: (in_ptr, in_ld, weight, bias, out_ptr,
        out_ld, rows, silu_in).  The device job table is built once per plan (all addresses are static)."""
        if not jobs:
            return
        arr = (LinearJob * len(jobs))()
        first = 0
        for j, (in_ptr, in_ld, w, b, out_ptr, out_ld, rows, silu) in zip(arr, jobs):
            n, k = w.shape
            j.in_, j.w, j.bias, j.out = in_ptr, w.data_ptr(), (b.data_ptr() if b is not None else None), out_ptr
            j.in_ld, j.out_ld, j.rows, j.k, j.n, j.silu_in, j.first_block = in_ld, out_ld, rows, k, n, silu, first
            first += int(self.lib.ldmb200_small_linear_blocks(rows, n))
        table = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(self.model._device())
        self.keep.append(table)
        self.emit("ldmb200_small_linear_batched", table.data_ptr(), len(jobs), first, what=what)

    def _prepass_vectors(self, emb: _Act, ctx: _Act):
        """Everything that depends only on (emb, context) and not on x, for the whole network, in three launches:
        all ResBlock emb projections (openaimodel.py:264) and — for a length-1 context, where softmax over one key
        is 1 and attn2(x) == to_out(to_v(ctx)) for every token — all cross-attention outputs."""
        m, B, L = self.model, self.B, self.L
        res = [(n, mod) for n, mod in m.named_modules() if isinstance(mod, ResBlock)]
        tbs = [(n, mod) for n, mod in m.named_modules() if isinstance(mod, BasicTransformerBlock)]
        self.vec: Dict[int, Tuple[int, int]] = {}   # id(module) -> (pointer, leading dim) of its per-sample row vector
        tot = sum(mod.out_channels for _, mod in res)
        ev = self.new(B, tot, F32)
        self.pinned.add(id(ev))
        jobs, off = [], 0
        for _, mod in res:
            lin = mod.emb_layers[1]
            jobs.append((emb.ptr, emb.c, lin.weight, lin.bias, ev.ptr + 4 * off, tot, B, 1))
            self.vec[id(mod)] = (ev.ptr + 4 * off, tot)
            off += mod.out_channels
        self.small_linear_batched(jobs, f"emb_layers x{len(jobs)}")
        tot = sum(mod.attn2.to_v.weight.shape[0] for _, mod in tbs)
        rows = B * max(L, 1)
        vall = self.new(rows, tot, F32)
        self.pinned.add(id(vall))
        jobs, off = [], 0
        for _, mod in tbs:
            w = mod.attn2.to_v.weight
            jobs.append((ctx.ptr, ctx.c, w, None, vall.ptr + 4 * off, tot, rows, 0))
            self.vec[("v", id(mod))] = (vall.ptr + 4 * off, tot)
            off += w.shape[0]
        self.small_linear_batched(jobs, f"attn2.to_v x{len(jobs)}")
        if L == 1:
            cv = self.new(B, tot, F32)
            self.pinned.add(id(cv))
            jobs, off = [], 0
            for _, mod in tbs:
                lo = mod.attn2.to_out[0]
                vp, vld = self.vec[("v", id(mod))]
                jobs.append((vp, vld, lo.weight, lo.bias, cv.ptr + 4 * off, tot, B, 0))
                self.vec[id(mod)] = (cv.ptr + 4 * off, tot)
                off += lo.weight.shape[0]
            self.small_linear_batched(jobs, f"attn2.to_out x{len(jobs)}")
        else:
            kall = self.new(rows, tot, F32)
            self.pinned.add(id(kall))
            jobs, off = [], 0
            for _, mod in tbs:
                w = mod.attn2.to_k.weight
                jobs.append((ctx.ptr, ctx.c, w, None, kall.ptr + 4 * off, tot, rows, 0))
                self.vec[("k", id(mod))] = (kall.ptr + 4 * off, tot)
                off += w.shape[0]
            self.small_linear_batched(jobs, f"attn2.to_k x{len(jobs)}")

    def groupnorm(self, x1: _Act, x2: Optional[_Act], gn: nn.GroupNorm, hw: int, silu: int, want_raw: bool):
        c = x1.c + (x2.c if x2 is not None else 0)
        y = self.new(x1.rows, c, self.act)
        raw = self.new(x1.rows, c, self.act) if want_raw else None
        s1, s2 = x1.stats, (x2.stats if x2 is not None else None)
        if s1 is not None and (x2 is None or s2 is not None):
            # statistics were taken by the GEMM epilogues that wrote x1 / x2: finalize + apply, x is read once
            self.emit("ldmb200_groupnorm_fused", x1.ptr, x1.c, s1.ptr, x2.ptr if x2 is not None else None,
                      x2.c if x2 is not None else 0, s2.ptr if s2 is not None else None, gn.weight.data_ptr(),
                      gn.bias.data_ptr(), y.ptr, raw.ptr if raw is not None else None, self.act, self.B, hw,
                      gn.num_groups, float(gn.eps), silu, self.gn_ws.data_ptr(),
                      bytes_=x1.rows * c * (4 + 2 * (2 if want_raw else 1)),
                      what=f"rows{x1.rows} C{c} fused" + (" cat" if x2 is not None else "") + (" raw" if want_raw else ""),
                      outs=(y, raw))
            return y, raw
        self.emit("ldmb200_groupnorm", x1.ptr, x1.c, x2.ptr if x2 is not None else None,
                  x2.c if x2 is not None else 0, gn.weight.data_ptr(), gn.bias.data_ptr(), y.ptr,
                  raw.ptr if raw is not None else None, self.act, self.B, hw, gn.num_groups, float(gn.eps), silu,
                  self.gn_ws.data_ptr(),
                  bytes_=x1.rows * c * (4 + (2 if self.tc else 4) * (2 if want_raw else 1)),
                  what=f"rows{x1.rows} C{c}" + (" cat" if x2 is not None else "") + (" raw" if want_raw else ""),
                  outs=(y, raw))
        return y, raw

    def layernorm(self, x: _Act, ln: nn.LayerNorm) -> _Act:
        y = self.new(x.rows, x.c, self.act)
        self.emit("ldmb200_layernorm", x.ptr, ln.weight.data_ptr(), ln.bias.data_ptr(), y.ptr, self.act, x.rows, x.c,
                  float(ln.eps), bytes_=x.rows * x.c * (4 + (2 if self.tc else 4)), what=f"rows{x.rows} C{x.c}", outs=(y,))
        return y

    def gemm(self, a: _Act, key: str, weight: torch.Tensor, bias: Optional[torch.Tensor], h_in: int, w_in: int,
             ksize: int = 1, stride: int = 1, upsample: int = 0, rowvec: Optional[Tuple[int, int]] = None,
             residual: Optional[_Act] = None, geglu: bool = False, out_dtype: int = F32,
             out: Optional[Tuple[int, int]] = None, out_nchw_ptr: Optional[int] = None,
             force_simt: bool = False, n_true: Optional[int] = None, want_stats: bool = False) -> Optional[_Act]:
        """Emits one (implicit) GEMM: ``ldmb200_igemm_tc`` in bf16 mode, ``ldmb200_igemm_simt`` otherwise.
        ``key`` names the packed weight; ``out=(ptr, ld)`` writes into an existing matrix (fused QKV in fp32
        mode); ``out_nchw_ptr`` selects the fp32 NCHW store of ``UNetModel.out``."""
        tc = self.tc and not force_simt
        d = GemmDesc()
        d.a, d.a2, d.c1, d.c2 = a.ptr, None, a.c, 0
        d.a_dtype = a.dtype
        d.batch, d.h_in, d.w_in = self.B, h_in, w_in
        d.ksize, d.stride, d.upsample = ksize, stride, upsample
        n_true = weight.shape[0] if n_true is None else n_true
        if tc:
            d.w = self.pack.ptr(key)
            d.n = self.pack.entries[key][1]
            assert a.c == self.pack.entries[key][2], (key, a.c, self.pack.entries[key])
            d.bias = (self.pack.fptr(key + ".bias") if geglu else (bias.data_ptr() if bias is not None else None))
        else:
            d.w = weight.data_ptr()
            d.n = n_true
            d.bias = bias.data_ptr() if bias is not None else None
        d.rowvec, d.rowvec_ld = rowvec if rowvec is not None else (None, 0)
        d.residual, d.residual_ld = (residual.ptr, residual.c) if residual is not None else (None, 0)
        d.geglu = 1 if geglu else 0
        ws = getattr(self, "splitk_ws", None)   # (the first-stage plans of vqdecode.py carry none)
        if tc and ws is not None and not geglu:
            d.splitk_ws, d.splitk_ws_bytes = ws.data_ptr(), ws.numel() * 4
        n_cols = n_true // 2 if geglu else n_true
        h_eff, w_eff = (2 * h_in, 2 * w_in) if upsample else (h_in, w_in)
        h_out, w_out = (h_eff // 2, w_eff // 2) if stride == 2 else (h_eff, w_eff)
        rows = self.B * h_out * w_out
        res = None
        if out_nchw_ptr is not None:
            d.out, d.out_dtype, d.out_ld, d.out_nchw, d.n_valid = out_nchw_ptr, F32, 0, 1, n_true
        elif out is not None:
            d.out, d.out_dtype, d.out_ld, d.out_nchw, d.n_valid = out[0], out_dtype, out[1], 0, n_cols
        else:
            res = self.new(rows, n_cols, out_dtype)
            d.out, d.out_dtype, d.out_ld, d.out_nchw, d.n_valid = res.ptr, out_dtype, n_cols, 0, n_cols
            if (want_stats and tc and out_dtype == F32 and not geglu and (h_out * w_out) % 32 == 0
                    and n_true % 32 == 0 and d.n == n_true):
                # GroupNorm statistics of this tensor, taken in the epilogue (consumed by groupnorm())
                res.stats = self.arena.alloc((rows // 32) * n_true * 2 * 4)
                d.colstats = res.stats.ptr
        self.keep.append(d)
        if tc and hasattr(self, "tc_gemms"):
            self.tc_gemms.append((d, getattr(self.pack, "nbytes", {}).get(key, 0), rows))   # (the training pack records no sizes: no prefetch)
        k_true = weight.shape[1] * ksize * ksize if weight.dim() >= 2 and n_true == weight.shape[0] else a.c
        # (folded upsample: `flops` stays the ALGORITHMIC count of the reference's op list, 2 M N 9C — SURVEY 8d — although
        # the four 2x2 convolutions execute 4/9 of it; bench.py reports the executed rate next to the algorithmic one)
        eb = lambda dt: 4 if dt == F32 else 2
        # algorithmic bytes: the input map once, the weights once, the output once (+ the fp32 residual it adds)
        abytes = (a.rows * a.c * eb(a.dtype) + n_true * k_true * (2 if tc else 4)
                  + rows * n_cols * (4 if out_nchw_ptr is not None else eb(out_dtype))
                  + (rows * n_cols * 4 if residual is not None else 0))
        self.emit("ldmb200_igemm_tc" if tc else "ldmb200_igemm_simt", C.byref(d), flops=2.0 * rows * n_true * k_true,
                  bytes_=abytes,
                  what=f"M{rows} N{n_true} K{k_true} ks{ksize} s{stride}" + (" geglu" if geglu else "") + (" upfold" if upsample and tc else "")
                  + (" res" if residual is not None else "") + ({BF16: " bf16out", F16: " f16out"}.get(out_dtype, "")),
                  outs=(res,))
        return res

    # ---- blocks ------------------------------------------------------------------------
    def resblock(self, m: ResBlock, name: str, x1: _Act, x2: Optional[_Act], H: int, W: int, emb: _Act) -> _Act:
        """openaimodel.py:255-275: GN-SiLU-conv, +emb, GN-SiLU-conv, skip(x)+h."""
        has_skip = isinstance(m.skip_connection, nn.Conv2d)
        if x2 is not None and not has_skip:
            raise NotImplementedError("identity skip over a concatenated input is not reachable from UNetModel")
        y1, raw = self.groupnorm(x1, x2, m.in_layers[0], H * W, 1, has_skip)
        c0 = m.in_layers[2]
        h1 = self.gemm(y1, name + ".in_layers.2", c0.weight, c0.bias, H, W, ksize=3, rowvec=self.vec[id(m)],
                       want_stats=True)
        self.free(y1)
        y2, _ = self.groupnorm(h1, None, m.out_layers[0], H * W, 1, False)
        self.free(h1)
        if has_skip:
            sc = m.skip_connection
            res = self.gemm(raw, name + ".skip_connection", sc.weight, sc.bias, H, W, ksize=1)
            self.free(raw)
        else:
            res = x1
        c1 = m.out_layers[3]
        h2 = self.gemm(y2, name + ".out_layers.3", c1.weight, c1.bias, H, W, ksize=3, residual=res, want_stats=True)
        self.free(y2)
        if has_skip:
            self.free(res)
        return h2

    def transformer_block(self, m: BasicTransformerBlock, name: str, t0: _Act, hw: int, ctx: _Act,
                          last: bool) -> _Act:
        """attention.py:211-215: x += attn1(LN1 x); x += attn2(LN2 x, ctx); x += ff(LN3 x)."""
        Cc, B, L = t0.c, self.B, self.L
        heads, dh = m.attn1.heads, m.attn1.dim_head
        if dh != 32:
            raise NotImplementedError(f"attention kernels are built for d_head = 32 (got {dh})")
        scale = float(m.attn1.scale)
        a = self.layernorm(t0, m.norm1)
        qkv = self.new(t0.rows, 3 * Cc, self.qkv_dt)
        if self.tc:
            self.gemm(a, name + ".attn1.qkv", m.attn1.to_q.weight, None, hw, 1, out_dtype=BF16,
                      out=(qkv.ptr, 3 * Cc), n_true=3 * Cc)
        else:
            for i, lin in enumerate((m.attn1.to_q, m.attn1.to_k, m.attn1.to_v)):
                self.gemm(a, "", lin.weight, None, hw, 1, out=(qkv.ptr + 4 * i * Cc, 3 * Cc))
        self.free(a)
        o = self.new(t0.rows, Cc, self.act)
        if self.tc:
            self.emit("ldmb200_attention_tc", qkv.ptr, 3 * Cc, 0, Cc, 2 * Cc, o.ptr, Cc, self.act, B, heads, hw, scale,
                      flops=4.0 * B * heads * hw * hw * dh, bytes_=t0.rows * Cc * 2 * 4, what=f"B{B} h{heads} N{hw}",
                      outs=(o,))
        else:
            self.emit("ldmb200_attention_simt", qkv.ptr, 3 * Cc, 0, qkv.ptr, Cc, qkv.ptr, 2 * Cc, 3 * Cc, o.ptr, Cc, B,
                      heads, hw, hw, scale, flops=4.0 * B * heads * hw * hw * dh)
        self.free(qkv)
        lo = m.attn1.to_out[0]
        a2m = m.attn2
        if L == 1:
            # one key: softmax == 1, attn2(x) == to_out(to_v(ctx)) for every token -> a per-sample
            # row vector added in the epilogue of attn1's output projection
            t2 = self.gemm(o, name + ".attn1.to_out.0", lo.weight, lo.bias, hw, 1, rowvec=self.vec[id(m)],
                           residual=t0)
            self.free(o, t0)
        else:
            t1 = self.gemm(o, name + ".attn1.to_out.0", lo.weight, lo.bias, hw, 1, residual=t0)
            self.free(o, t0)
            a2 = self.layernorm(t1, m.norm2)
            q = self.gemm(a2, name + ".attn2.to_q", a2m.to_q.weight, None, hw, 1)
            self.free(a2)
            (kp, kld), (vp, _) = self.vec[("k", id(m))], self.vec[("v", id(m))]
            o2 = self.new(t1.rows, Cc, F32)
            self.emit("ldmb200_attention_simt", q.ptr, Cc, 0, kp, 0, vp, 0, kld, o2.ptr, Cc, B, a2m.heads, hw, L,
                      float(a2m.scale), flops=4.0 * B * a2m.heads * hw * L * dh)
            self.free(q)
            lo2 = a2m.to_out[0]
            t2 = self.gemm(o2, "", lo2.weight, lo2.bias, hw, 1, residual=t1, force_simt=True)
            self.free(o2, t1)
        f = self.layernorm(t2, m.norm3)
        gp = m.ff.net[0].proj
        g = self.gemm(f, name + ".ff.net.0.proj", gp.weight, gp.bias, hw, 1, geglu=True, out_dtype=self.act)
        self.free(f)
        l2 = m.ff.net[2]
        t3 = self.gemm(g, name + ".ff.net.2", l2.weight, l2.bias, hw, 1, residual=t2,
                       out_dtype=self.act if last else F32)
        self.free(g, t2)
        return t3

    def spatial_transformer(self, m: SpatialTransformer, name: str, x: _Act, H: int, W: int, ctx: _Act) -> _Act:
        """attention.py:250-261: x_in + proj_out(blocks(proj_in(GN(x))))."""
        y, _ = self.groupnorm(x, None, m.norm, H * W, 0, False)
        t = self.gemm(y, name + ".proj_in", m.proj_in.weight, m.proj_in.bias, H * W, 1)
        self.free(y)
        nb = len(m.transformer_blocks)
        for i, blk in enumerate(m.transformer_blocks):
            t = self.transformer_block(blk, f"{name}.transformer_blocks.{i}", t, H * W, ctx, i == nb - 1)
        out = self.gemm(t, name + ".proj_out", m.proj_out.weight, m.proj_out.bias, H * W, 1, residual=x,
                        want_stats=True)
        self.free(t)
        return out

    def downsample(self, m: Downsample, name: str, x: _Act, H: int, W: int) -> _Act:
        """openaimodel.py:151-160: conv 3x3 stride 2 pad 1."""
        if self.tc:
            s = self.new(x.rows, x.c, self.act)
            self.emit("ldmb200_space_to_depth2", x.ptr, s.ptr, self.act, self.B, H, W, x.c, outs=(s,))
            out = self.gemm(s, name + ".op", m.op.weight, m.op.bias, H, W, ksize=3, stride=2, want_stats=True)
            self.free(s)
            return out
        return self.gemm(x, "", m.op.weight, m.op.bias, H, W, ksize=3, stride=2)

    def upsample(self, m: Upsample, name: str, x: _Act, H: int, W: int) -> _Act:
        """openaimodel.py:109-119: nearest 2x, then conv 3x3."""
        if self.tc and self.fold_upsample and H & (H - 1) == 0 and W & (W - 1) == 0:
            # K10 folded into the conv: the 16-bit copy of the LOW-RES map is the operand of four 2x2 convolutions (one per
            # output parity, tap-summed weights): 4/9 of the flops, no materialised 4x larger map
            lo = self.new(x.rows, x.c, self.act)
            self.emit("ldmb200_cast_rows", x.ptr, lo.ptr, self.act, x.rows * x.c, bytes_=x.rows * x.c * 6, outs=(lo,))
            out = self.gemm(lo, name + ".conv.fold", m.conv.weight, m.conv.bias, H, W, ksize=3, upsample=1,
                            want_stats=(H * W) % 32 == 0)
            self.free(lo)
            return out
        if self.tc:
            u = self.new(4 * x.rows, x.c, self.act)
            self.emit("ldmb200_upsample2x", x.ptr, u.ptr, self.act, self.B, H, W, x.c, outs=(u,))
            out = self.gemm(u, name + ".conv", m.conv.weight, m.conv.bias, 2 * H, 2 * W, ksize=3, want_stats=True)
            self.free(u)
            return out
        return self.gemm(x, "", m.conv.weight, m.conv.bias, H, W, ksize=3, upsample=1)

    # ---- whole network -------------------------------------------------------------------
    def run_block(self, block: nn.Sequential, name: str, h: _Act, x2: Optional[_Act], H: int, W: int, emb: _Act,
                  ctx: _Act):
        """TimestepEmbedSequential.forward (openaimodel.py:80-88) over parameter holders.  ``x2`` is the
        skip tensor virtually concatenated to ``h`` for the block's first ResBlock (openaimodel.py:736)."""
        for j, layer in enumerate(block):
            lname = f"{name}.{j}"
            if isinstance(layer, ResBlock):
                new = self.resblock(layer, lname, h, x2, H, W, emb)
                self.free(h, x2)
                x2 = None
            elif isinstance(layer, SpatialTransformer):
                new = self.spatial_transformer(layer, lname, h, H, W, ctx)
                self.free(h)
            elif isinstance(layer, Downsample):
                new = self.downsample(layer, lname, h, H, W)
                self.free(h)
                H, W = H // 2, W // 2
            elif isinstance(layer, Upsample):
                new = self.upsample(layer, lname, h, H, W)
                self.free(h)
                H, W = 2 * H, 2 * W
            elif isinstance(layer, nn.Conv2d):
                new = self.gemm(h, lname, layer.weight, layer.bias, H, W, ksize=3, want_stats=True)
                self.free(h)
            else:
                raise NotImplementedError(f"{lname}: {type(layer).__name__}")
            h = new
        assert x2 is None
        return h, H, W

    def _build(self):
        m, B, H, W = self.model, self.B, self.H, self.W
        self.pinned = set()
        mc, ted = m.model_channels, 4 * m.model_channels
        # --- t_emb -> emb (util.py:151-171, openaimodel.py:723-728)
        temb = self.new(B, mc, F32)
        self.emit("ldmb200_timestep_embedding", self.t_in.data_ptr(), temb.ptr, B, mc, 10000.0)
        e1 = self.small_linear(temb.ptr, mc, m.time_embed[0], B, 0)
        emb = self.small_linear(e1.ptr, ted, m.time_embed[2], B, 1)
        if m.num_classes is not None:
            emb_t, y_in, table = self.view_f32(emb), self.y_in, m.label_emb
            self.py_ops[len(self.ops)] = lambda: emb_t.add_(table.weight.detach()[y_in])
        self.pinned.add(id(emb))
        ctx = _Act(_Buf(self.ctx_in.view(torch.uint8).view(-1)), B * max(self.L, 1), self.ctx_in.shape[-1], F32)
        self.pinned.add(id(ctx))
        self._prepass_vectors(emb, ctx)
        # --- input layout: NCHW fp32 -> channels-last (zero padded to 32 channels on the tensor-core path)
        cin = m.in_channels
        cpad = _tc_cpad(cin) if self.tc else cin
        h = self.new(B * H * W, cpad, self.act)
        self.emit("ldmb200_nchw_to_nhwc", self.x_in.data_ptr(), h.ptr, B, cin, H * W, cpad, self.act)
        hs: List[Tuple[_Act, int, int]] = []
        for i, blk in enumerate(m.input_blocks):
            h, H, W = self.run_block(blk, f"input_blocks.{i}", h, None, H, W, emb, ctx)
            hs.append(h)
            self.pinned.add(id(h))
        h, H, W = self.run_block(m.middle_block, "middle_block", h, None, H, W, emb, ctx)
        for i, blk in enumerate(m.output_blocks):
            skip = hs.pop()
            self.pinned.discard(id(skip))
            if i == 0:
                pass  # h is the middle block's output (not pinned); the last input-block output is `skip`
            h, H, W = self.run_block(blk, f"output_blocks.{i}", h, skip, H, W, emb, ctx)
        y, _ = self.groupnorm(h, None, m.out[0], H * W, 1, False)
        self.free(h)
        oc = m.out[2]
        self.gemm(y, "out.2", oc.weight, oc.bias, H, W, ksize=3, out_nchw_ptr=self.out.data_ptr())
        self.free(y)

    def kernel_launches(self) -> int:
        """CUDA kernels one replay of the plan launches (a C-ABI call is one kernel, except: GroupNorm = statistics /
        finalize + apply, the folded Upsample convolution = four parity launches)."""
        n = 0
        for name, _fl, _by, what in self.meta:
            n += 2 if name.startswith("ldmb200_groupnorm") else 4 if "upfold" in what else 1
        return n

    def run_timed(self, stream_obj) -> List[tuple]:
        """Runs the plan once with a CUDA event pair around every launch on ``stream_obj`` (the launching stream);
        returns ``[(kernel name, ms, algorithmic flops, algorithmic bytes)]``.  Measurement aid for bench.py."""
        evs = []
        for i, (fn, args, name) in enumerate(self.ops):
            if i in self.py_ops:
                self.py_ops[i]()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream_obj)
            rc = fn(*args, stream_obj.cuda_stream)
            e1.record(stream_obj)
            if rc:
                _cabi.check(rc, name)
            evs.append((e0, e1))
        stream_obj.synchronize()
        return [(m[0], e0.elapsed_time(e1), m[1], m[2], m[3]) for m, (e0, e1) in zip(self.meta, evs)]

    def run_checked(self, stream_obj) -> Dict[str, int]:
        """Debug run: after every launch, counts the elements of its fp16 outputs that sit at the clamp of the saturating
        conversion (|v| == 65504) or are not finite.  Returns ``{"<op index> <kernel> <shape>": count}`` for the
        launches that saturated (empty = the fp16 operand range was never exceeded)."""
        counts = torch.zeros(len(self.ops), dtype=torch.int64, device=self.model._device())
        for i, (fn, args, name) in enumerate(self.ops):
            if i in self.py_ops:
                self.py_ops[i]()
            rc = fn(*args, stream_obj.cuda_stream)
            if rc:
                _cabi.check(rc, name)
            for buf, n, dt in self.op_outs.get(i, ()):
                if dt == F16:   # one counting launch per fp16 output, no host sync inside the eval
                    _cabi.call("ldmb200_count_f16_saturated", buf.ptr, n, counts.data_ptr() + 8 * i, stream_obj.cuda_stream)
        self.runs += 1
        host = counts.cpu().tolist()
        return {f"{i} {self.ops[i][2][8:]} {self.meta[i][3]}": c for i, c in enumerate(host) if c}

    def run(self, stream: int):
        py = self.py_ops
        for i, (fn, args, name) in enumerate(self.ops):
            if py and i in py:
                py[i]()
            rc = fn(*args, stream)
            if rc:
                _cabi.check(rc, name)
        self.runs += 1


# ----------------------------------------------------------------------------------------------
# the drop-in module
# ----------------------------------------------------------------------------------------------

class UNetModel(nn.Module):
    """Drop-in for the reference ``UNetModel`` (openaimodel.py:413-742); see the module docstring.

    Supported configuration = what every shipped YAML uses: ``dims=2``,
    ``use_spatial_transformer=True``, ``conv_resample=True``, ``resblock_updown=False``,
    ``use_scale_shift_norm=False``, ``n_embed=None``, ``num_head_channels=32`` (or ``num_heads`` giving
    32-wide heads).  Anything else raises ``NotImplementedError`` at construction — never a silent fallback.
    """

    MAX_PLANS = 4

    def __init__(self, image_size, in_channels, model_channels, out_channels, num_res_blocks,
                 attention_resolutions, dropout=0, channel_mult=(1, 2, 4, 8), conv_resample=True, dims=2,
                 num_classes=None, use_checkpoint=False, use_fp16=False, num_heads=-1, num_head_channels=-1,
                 num_heads_upsample=-1, use_scale_shift_norm=False, resblock_updown=False,
                 use_new_attention_order=False, use_spatial_transformer=False, transformer_depth=1,
                 context_dim=None, n_embed=None, legacy=True):
        super().__init__()
        if use_spatial_transformer:
            assert context_dim is not None, "use_spatial_transformer needs context_dim (openaimodel.py:471-472)"
        if context_dim is not None:
            assert use_spatial_transformer, "context_dim needs use_spatial_transformer (openaimodel.py:474-475)"
        attention_resolutions = _aslist(attention_resolutions)
        channel_mult = _aslist(channel_mult)
        context_dim = _aslist(context_dim)
        if isinstance(context_dim, list):
            if len(context_dim) != 1:
                raise NotImplementedError("a list-valued context_dim with more than one entry")
            context_dim = int(context_dim[0])
        for flag, ok, what in ((dims, 2, "dims"), (bool(conv_resample), True, "conv_resample"),
                               (bool(use_scale_shift_norm), False, "use_scale_shift_norm"),
                               (bool(resblock_updown), False, "resblock_updown"),
                               (bool(use_spatial_transformer), True, "use_spatial_transformer"),
                               (n_embed, None, "n_embed"), (bool(use_fp16), False, "use_fp16")):
            if flag != ok:
                raise NotImplementedError(f"UNetModel({what}={flag!r}) is outside the B200 hot path "
                                          f"(only {what}={ok!r} is used by the reference configs)")
        if float(dropout) != 0.0:
            raise NotImplementedError("dropout > 0 (all reference configs use 0)")
        if num_heads_upsample == -1:
            num_heads_upsample = num_heads
        if num_heads == -1:
            assert num_head_channels != -1, "Either num_heads or num_head_channels has to be set"
        if num_head_channels == -1:
            assert num_heads != -1, "Either num_heads or num_head_channels has to be set"

        self.image_size = image_size
        self.in_channels = in_channels
        self.model_channels = model_channels
        self.out_channels = out_channels
        self.num_res_blocks = num_res_blocks
        self.attention_resolutions = attention_resolutions
        self.dropout = dropout
        self.channel_mult = channel_mult
        self.conv_resample = conv_resample
        self.num_classes = num_classes
        self.use_checkpoint = use_checkpoint
        self.dtype = torch.float32
        self.num_heads = num_heads
        self.num_head_channels = num_head_channels
        self.num_heads_upsample = num_heads_upsample
        self.predict_codebook_ids = False
        self.context_dim = context_dim
        self.transformer_depth = transformer_depth

        def heads_for(ch):
            # openaimodel.py:541-549: with the spatial transformer dim_head always ends up ch // num_heads
            nh = num_heads if num_head_channels == -1 else ch // num_head_channels
            return nh, ch // nh

        def st(ch):
            nh, dh = heads_for(ch)
            return SpatialTransformer(ch, nh, dh, transformer_depth, context_dim)

        ted = model_channels * 4
        self.time_embed = nn.Sequential(nn.Linear(model_channels, ted), _Slot(), nn.Linear(ted, ted))
        if num_classes is not None:
            self.label_emb = nn.Embedding(num_classes, ted)
        self.input_blocks = nn.ModuleList(
            [TimestepEmbedSequential(nn.Conv2d(in_channels, model_channels, 3, padding=1))])
        chans = [model_channels]
        ch, ds = model_channels, 1
        for level, mult in enumerate(channel_mult):
            for _ in range(num_res_blocks):
                layers = [ResBlock(ch, ted, mult * model_channels)]
                ch = mult * model_channels
                if ds in attention_resolutions:
                    layers.append(st(ch))
                self.input_blocks.append(TimestepEmbedSequential(*layers))
                chans.append(ch)
            if level != len(channel_mult) - 1:
                self.input_blocks.append(TimestepEmbedSequential(Downsample(ch)))
                chans.append(ch)
                ds *= 2
        self.middle_block = TimestepEmbedSequential(ResBlock(ch, ted, ch), st(ch), ResBlock(ch, ted, ch))
        self.output_blocks = nn.ModuleList([])
        for level, mult in list(enumerate(channel_mult))[::-1]:
            for i in range(num_res_blocks + 1):
                ich = chans.pop()
                layers = [ResBlock(ch + ich, ted, model_channels * mult)]
                ch = model_channels * mult
                if ds in attention_resolutions:
                    layers.append(st(ch))
                if level and i == num_res_blocks:
                    layers.append(Upsample(ch))
                    ds //= 2
                self.output_blocks.append(TimestepEmbedSequential(*layers))
        self.out = nn.Sequential(nn.GroupNorm(32, ch), _Slot(),
                                 _zero(nn.Conv2d(model_channels, out_channels, 3, padding=1)))

        # ---- runtime state (not parameters / buffers: state_dict stays identical to the reference)
        self.precision = os.environ.get("LDMB200_PRECISION", "f16")
        # CUDA-graph replay of the static launch plan is the default (the benchmarked configuration); the first call of a
        # plan runs eagerly (lazy per-kernel setup), the second captures.  LDMB200_CUDA_GRAPH=0 launches eagerly.
        self.use_cuda_graph = os.environ.get("LDMB200_CUDA_GRAPH", "1") == "1"
        # fp16 operand range guard, evaluated on the first call of every plan: "warn" (default) = fall back to bf16
        # operands for the whole model with a RuntimeWarning when a launch saturates at +-65504, "raise", or "off"
        self.f16_guard = os.environ.get("LDMB200_F16_GUARD", "warn")
        # split-K of the tensor-core GEMMs whose grid would leave most SMs idle (batch 1-4 sampling).  Opt-in (split_k = True
        # or LDMB200_SPLITK=1): measured 2-3 % per eval at batch 1-2 (igemm_tc.cu), and it gives up the bit-identity of a
        # sample's eps BETWEEN batch sizes that take different split decisions (fp32 summation order; every batch shape stays
        # bit-reproducible run to run).  Off, a sample's eps is bit-identical under any sharding of the batch.
        self._split_k = os.environ.get("LDMB200_SPLITK", "0") == "1"
        self._plans: Dict[tuple, _Plan] = {}
        self._packs: Dict[int, _WeightPack] = {}   # operand dtype -> packed weights
        self._frozen = 0
        self._pack_valid: set = set()              # dtypes whose pack matches the current parameters
        self.launches = 0   # kernels launched by forward() so far (bench.py's gpu_launches)
        self._train_state = None  # dsml_thesis_b200.train._TrainState (gradient arena, training plans)
        self._grad_sync = None    # dsml_thesis_b200.train.GradAllReduce attached to this model (explicit or automatic)
        self._grad_bucket_hook = None

    # ---- reference API stubs (openaimodel.py:694-708): the fp16 conversion is a no-op there too -------
    def convert_to_fp16(self):
        pass

    def convert_to_fp32(self):
        pass

    # ---- runtime plumbing ------------------------------------------------------------------------
    def _device(self):
        return self.out[2].weight.device

    def _apply(self, fn, *a, **k):
        # .to()/.cuda()/.float() may move parameter storage: drop everything that holds raw pointers
        self._train_state = None
        self._grad_sync = None          # holds views of the old gradient arena
        self._grad_bucket_hook = None
        self._plans = {}
        self._packs = {}
        self._pack_valid = set()
        return super()._apply(fn, *a, **k)

    def _weight_pack(self, dtype: int) -> _WeightPack:
        if dtype not in self._packs:
            self._packs[dtype] = _WeightPack(self, dtype)
            self._pack_valid.discard(dtype)
        return self._packs[dtype]

    def invalidate_weight_cache(self):
        """Call after writing parameters in place through ``.data`` while the weights are frozen (``LitEma.copy_to``
        / ``restore``, ema.py:46-76, bump neither ``_version`` nor ``data_ptr``)."""
        self._pack_valid = set()

    @contextlib.contextmanager
    def frozen_weights(self):
        """Inside this context the bf16 weight pack is refreshed once (on the first forward) instead of on every
        forward.  ``DDIMSampler.sample`` wraps its loop in it; weights cannot change inside a sampling loop."""
        self._frozen += 1
        if self._frozen == 1:
            self._pack_valid = set()
        try:
            yield self
        finally:
            self._frozen -= 1

    @torch.no_grad()
    def f16_saturation_report(self, x, timesteps, context, y=None) -> Dict[str, int]:
        """One eager eval in ``"f16"`` mode that counts, launch by launch, the fp16 operand elements that hit the
        saturating conversion's clamp (±65504) — the failure mode fp16 operands have and bf16 operands do not.
        Synthetic weights never get near it; a trained checkpoint with outlier activations (the 16-bit copy of the
        residual stream after the last transformer block, GEGLU products) can.  Empty dict = safe for these inputs.
        ``f16_guard`` (below) runs this automatically on the first call of every plan."""
        B, _, H, W = x.shape
        with torch.cuda.device(x.device):
            plan = self._get_plan(B, H, W, int(context.shape[1]), "f16")
            plan.x_in.copy_(x)
            plan.t_in.copy_(timesteps)
            plan.ctx_in.copy_(context)
            if y is not None:
                plan.y_in.copy_(y)
            cur = torch.cuda.current_stream()
            plan.pack.run(cur.cuda_stream)
            self._pack_valid.add(plan.act)
            return plan.run_checked(cur)

    @property
    def split_k(self) -> bool:
        return self._split_k

    @split_k.setter
    def split_k(self, on: bool):
        if bool(on) != self._split_k:
            self._split_k = bool(on)
            self._plans.clear()   # the workspace pointer is part of the recorded launches

    def _get_plan(self, B, H, W, L, mode) -> _Plan:
        key = (B, H, W, L, mode)
        plan = self._plans.get(key)
        if plan is None:
            while len(self._plans) >= self.MAX_PLANS:
                self._plans.pop(next(iter(self._plans)))
            plan = _Plan(self, B, H, W, L, mode)
            self._plans[key] = plan
        return plan

    def forward(self, x, timesteps=None, context=None, y=None, **kwargs):
        """openaimodel.py:710-742.  x [N,C,H,W], timesteps int [N], context [N,L,context_dim] -> [N,out,H,W]."""
        assert (y is not None) == (self.num_classes is not None), \
            "must specify y if and only if the model is class-conditional"
        return self._forward(x, timesteps, context, y, clone=True)

    def _forward(self, x, timesteps, context, y=None, clone=True):
        if not x.is_cuda:
            raise _cabi.Ldmb200Error("dsml_thesis_b200.UNetModel runs on a CUDA device only (no CPU fallback)")
        if torch.is_grad_enabled() and self.training:
            # DDPM training step (ddpm.py:1014-1047): forward that keeps what the backward kernels need + autograd node
            from .train import unet_train_forward
            return unet_train_forward(self, x, timesteps, context, y)
        if torch.is_grad_enabled() and x.requires_grad:
            raise NotImplementedError("gradients w.r.t. the latent in eval mode (guidance by gradient) are not built; "
                                      "call under torch.no_grad(), or model.train() for the training step")
        if context is None:
            raise ValueError("context is required (use_spatial_transformer=True, attention.py:170-176)")
        if timesteps is None or timesteps.dtype.is_floating_point:
            raise ValueError("timesteps must be an integer tensor of shape [N]")
        mode = self.precision
        if mode not in ("f16", "bf16", "fp32"):
            raise ValueError(f"precision must be 'f16', 'bf16' or 'fp32', got {mode!r}")
        B, Cin, H, W = x.shape
        assert Cin == self.in_channels, f"expected {self.in_channels} input channels, got {Cin}"
        assert context.dim() == 3 and context.shape[0] == B and context.shape[2] == self.context_dim, \
            f"context must be [N, L, {self.context_dim}], got {tuple(context.shape)}"
        assert timesteps.shape == (B,), f"timesteps must be [N], got {tuple(timesteps.shape)}"
        down = 2 ** (len(self.channel_mult) - 1)
        assert H % down == 0 and W % down == 0, f"H, W must be divisible by {down}"
        with torch.cuda.device(x.device):
            plan = self._get_plan(B, H, W, int(context.shape[1]), mode)
            plan.x_in.copy_(x)
            plan.t_in.copy_(timesteps)
            plan.ctx_in.copy_(context)
            if y is not None:
                assert y.shape == (B,)
                plan.y_in.copy_(y)
            need_pack = plan.tc and not (self._frozen and plan.act in self._pack_valid)
            self.launches += plan.kernel_launches() + (1 if need_pack else 0)
            if mode == "f16" and plan.runs == 0 and self.f16_guard != "off":
                # first call of this plan (eager anyway): the checked run.  fp16 operands that saturate silently change
                # the result; bf16 operands cannot saturate — fall back to them for the whole model, loudly.
                cur = torch.cuda.current_stream()
                if need_pack:
                    plan.pack.run(cur.cuda_stream)
                    self._pack_valid.add(plan.act)
                bad = plan.run_checked(cur)
                if bad:
                    msg = ("dsml_thesis_b200.UNetModel: %d launch(es) saturated their fp16 operands at +-65504 (first: %s, "
                           "%d elements)" % (len(bad), next(iter(bad)), next(iter(bad.values()))))
                    if self.f16_guard == "raise":
                        raise _cabi.Ldmb200Error(msg + "; set model.precision = 'bf16'")
                    import warnings
                    warnings.warn(msg + "; switching this model to precision='bf16' (bf16 operands cannot saturate; "
                                  "eps error ~1.2e-2 instead of ~1.5e-3)", RuntimeWarning)
                    self.precision = "bf16"
                    return self._forward(x, timesteps, context, y, clone)
            elif self.use_cuda_graph and plan.runs >= 1:
                if plan.graph is None:
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        plan.run(torch.cuda.current_stream().cuda_stream)
                    plan.graph = g
                if need_pack:
                    plan.pack.run(torch.cuda.current_stream().cuda_stream)
                    self._pack_valid.add(plan.act)
                plan.graph.replay()
            else:
                stream = torch.cuda.current_stream().cuda_stream
                if need_pack:
                    plan.pack.run(stream)
                    self._pack_valid.add(plan.act)
                plan.run(stream)
            out = plan.out
            if out.dtype != x.dtype:
                return out.to(x.dtype)
            return out.clone() if clone else out
