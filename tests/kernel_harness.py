"""Thin helpers that call ONE C-ABI kernel on torch CUDA tensors and build the matching plain-PyTorch fp32
reference.  Shared by the ``-m gpu`` parity tests and ``tools/gpu_diag.py``.  Test infrastructure only."""
import ctypes as C
import math

import torch
import torch.nn.functional as F

from dsml_thesis_b200 import _cabi
from dsml_thesis_b200._cabi import BF16, F16, F32, GemmDesc, LinearJob, PackJob
from dsml_thesis_b200.unet import _tc_bn, _tc_cpad, _tc_npad

DEV = "cuda"


def stream():
    return torch.cuda.current_stream().cuda_stream


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def to_cl(x, dtype=torch.float32):
    """NCHW -> channels-last matrix [B*H*W, C]."""
    B, Cc, H, W = x.shape
    return x.permute(0, 2, 3, 1).reshape(B * H * W, Cc).contiguous().to(dtype)


def from_cl(m, B, H, W):
    return m.float().reshape(B, H, W, -1).permute(0, 3, 1, 2).contiguous()


def ptr(t):
    return t.data_ptr() if t is not None else None


TORCH_DT = {F32: torch.float32, BF16: torch.bfloat16, F16: torch.float16}
CODE_DT = {v: k for k, v in TORCH_DT.items()}


def rnd16(x, dt):
    """fp32 -> rounded through the 16-bit operand type dt -> fp32 (what the tensor core sees)."""
    return x.to(TORCH_DT[dt]).float()


# ---------------------------------------------------------------------------------------------
def pack_weight(w, kind=0, bias=None, parts=None, dtype=BF16):
    """fp32 [n, c, k, k] (or [n, c]) -> packed bf16 [n_pad, k*k, c_pad] through ldmb200_pack_weights.
    Returns (packed tensor, packed fp32 bias or None, n_pad, c_pad)."""
    parts = parts or [w]
    ws = [p.reshape(p.shape[0], p.shape[1], *(p.shape[2:] if p.dim() == 4 else (1, 1))).contiguous() for p in parts]
    ks = ws[0].shape[2]
    c = ws[0].shape[1]
    n_tot = sum(p.shape[0] for p in ws)
    n_pad, c_pad = _tc_npad(n_tot), _tc_cpad(c)
    if kind == 4:   # folded Upsample conv: four tap-summed 2x2 weight sets [4][n_pad][4][c_pad]
        dst = torch.full((4, n_pad, 4, c_pad), float("nan"), dtype=TORCH_DT[dtype], device=DEV)
        arr = (PackJob * 1)()
        j = arr[0]
        (j.src, j.dst, j.kind, j.n, j.c, j.ksize, j.n_pad, j.c_pad, j.row_offset, j.block_n) = (
            ws[0].data_ptr(), dst.data_ptr(), 4, n_tot, c, 3, n_pad, c_pad, 0, _tc_bn(n_pad))
        j.dst_dtype = dtype
        dev_jobs = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(DEV)
        _cabi.call("ldmb200_pack_weights", dev_jobs.data_ptr(), 1, stream())
        torch.cuda.synchronize()
        return dst, None, n_pad, c_pad
    dst = torch.full((n_pad, ks * ks, c_pad), float("nan"), dtype=TORCH_DT[dtype], device=DEV)
    fb = torch.full((n_tot,), float("nan"), dtype=torch.float32, device=DEV) if (kind == 1 and bias is not None) else None
    jobs = []
    row = 0
    for i, p in enumerate(ws):
        rows = (n_pad - row) if i == len(ws) - 1 else p.shape[0]
        jobs.append((p.data_ptr(), dst.data_ptr(), kind, p.shape[0], c, ks, rows, c_pad, row, _tc_bn(n_pad, kind == 1)))
        row += p.shape[0]
    if fb is not None:
        jobs.append((bias.data_ptr(), fb.data_ptr(), 2, n_tot, 1, 1, n_tot, 1, 0, _tc_bn(n_pad, True)))
    arr = (PackJob * len(jobs))()
    for j, v in zip(arr, jobs):
        (j.src, j.dst, j.kind, j.n, j.c, j.ksize, j.n_pad, j.c_pad, j.row_offset, j.block_n) = v
        j.dst_dtype = dtype
    dev_jobs = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(DEV)
    _cabi.call("ldmb200_pack_weights", dev_jobs.data_ptr(), len(jobs), stream())
    torch.cuda.synchronize()
    return dst, fb, n_pad, c_pad


def geglu_perm(n_total, block_n):
    """Source row of every packed row (mirror of geglu_src_row in elementwise.cu)."""
    idx = []
    for rp in range(n_total):
        t, rem = divmod(rp, block_n)
        qd, i = rem >> 5, rem & 31
        col = t * (block_n >> 1) + qd * 16 + (i & 15)
        idx.append(col if i < 16 else (n_total >> 1) + col)
    return torch.tensor(idx, device=DEV)


def igemm(mode, a, B, h_in, w_in, w, bias=None, ksize=1, stride=1, upsample=0, a2=None, rowvec=None,
          residual=None, geglu=False, out_dtype=F32, out_nchw=False, colstats=None, splitk_ws=None):
    """mode 'simt' | 'tc'.  a / a2: channels-last matrices [B*h*w, c] (fp32 for simt, bf16 for tc; for tc+stride 2
    ``a`` is the space-to-depth layout).  w: fp32 [n, c1+c2, k, k].  Returns the output tensor."""
    d = GemmDesc()
    keep = []
    n_true = w.shape[0]
    d.a, d.a2 = a.data_ptr(), ptr(a2)
    d.c1, d.c2 = a.shape[1], (a2.shape[1] if a2 is not None else 0)
    d.a_dtype = CODE_DT[a.dtype]
    d.batch, d.h_in, d.w_in = B, h_in, w_in
    d.ksize, d.stride, d.upsample = ksize, stride, upsample
    if mode == "tc":
        wp, fb, n_pad, c_pad = pack_weight(w, kind=4 if upsample else (1 if geglu else 0), bias=bias, dtype=d.a_dtype)
        assert c_pad == d.c1 + d.c2 or d.c2 == 0, "test harness: concat inputs must be multiples of 32"
        keep += [wp, fb]
        d.w, d.n = wp.data_ptr(), n_pad
        d.bias = ptr(fb) if geglu else ptr(bias)
    else:
        wc = w.contiguous()
        keep.append(wc)
        d.w, d.n, d.bias = wc.data_ptr(), n_true, ptr(bias)
    d.rowvec, d.rowvec_ld = ptr(rowvec), (rowvec.shape[1] if rowvec is not None else 0)
    d.residual, d.residual_ld = ptr(residual), (residual.shape[1] if residual is not None else 0)
    d.geglu = int(geglu)
    n_cols = n_true // 2 if geglu else n_true
    h_eff, w_eff = (2 * h_in, 2 * w_in) if upsample else (h_in, w_in)
    h_out, w_out = (h_eff // 2, w_eff // 2) if stride == 2 else (h_eff, w_eff)
    M = B * h_out * w_out
    tdt = TORCH_DT[out_dtype]
    if out_nchw:
        out = torch.full((B, n_true, h_out, w_out), float("nan"), dtype=torch.float32, device=DEV)
        d.out, d.out_dtype, d.out_ld, d.out_nchw, d.n_valid = out.data_ptr(), F32, 0, 1, n_true
    else:
        # 64 guard rows behind the output: a store past row M (ragged last tile) would show up in them
        whole = torch.full((M + 64, n_cols), float("nan"), dtype=tdt, device=DEV)
        whole[M:] = 7.0
        igemm.guard = whole[M:]
        out = whole[:M]
        d.out, d.out_dtype, d.out_ld, d.out_nchw, d.n_valid = out.data_ptr(), out_dtype, n_cols, 0, n_cols
    if colstats is not None:
        d.colstats = colstats.data_ptr()
    if splitk_ws is not None:   # float32 tensor: LDMB200_SPLITK_COUNTER_BYTES of zeroed counters + partials
        d.splitk_ws, d.splitk_ws_bytes = splitk_ws.data_ptr(), splitk_ws.numel() * 4
    _cabi.call("ldmb200_igemm_tc" if mode == "tc" else "ldmb200_igemm_simt", C.byref(d), stream())
    torch.cuda.synchronize()
    if not out_nchw:
        assert bool((igemm.guard == 7.0).all()), "igemm wrote past the last output row"
    return out


def igemm_reference(x_nchw, w, bias=None, stride=1, upsample=0, rowvec=None, residual_nchw=None, geglu=False):
    """fp32 torch reference on NCHW tensors; returns NCHW."""
    x = x_nchw.float()
    if upsample:
        x = F.interpolate(x, scale_factor=2, mode="nearest")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    y = F.conv2d(x.double(), w.double(), bias.double() if bias is not None else None, stride=stride,
                 padding=w.shape[2] // 2).float()
    if geglu:
        a, g = y.chunk(2, dim=1)
        y = a * F.gelu(g)
    if rowvec is not None:
        y = y + rowvec[:, :, None, None]
    if residual_nchw is not None:
        y = y + residual_nchw
    return y


def space_to_depth(x_nchw, dtype):
    """NCHW -> the [B,4,H/2,W/2,C] plane layout as a matrix [B*4*H/2*W/2, C] via the kernel."""
    B, Cc, H, W = x_nchw.shape
    src = to_cl(x_nchw)
    out = torch.empty(B * H * W, Cc, dtype=dtype, device=DEV)
    _cabi.call("ldmb200_space_to_depth2", src.data_ptr(), out.data_ptr(), CODE_DT[dtype], B, H, W, Cc, stream())
    return out


def attention(mode, q, k, v, heads, scale, out_dtype=BF16):
    """q [B,Nq,C], k/v [B,Nkv,C] fp32 -> out [B,Nq,C].  tc: self-attention layout (fused QKV, bf16)."""
    B, Nq, Cc = q.shape
    Nkv = k.shape[1]
    if mode == "tc":
        assert Nq == Nkv
        qkv = torch.cat([q, k, v], dim=2).reshape(B * Nq, 3 * Cc).to(torch.bfloat16).contiguous()
        out = torch.full((B * Nq, Cc), float("nan"), dtype=TORCH_DT[out_dtype], device=DEV)
        _cabi.call("ldmb200_attention_tc", qkv.data_ptr(), 3 * Cc, 0, Cc, 2 * Cc, out.data_ptr(), Cc, out_dtype, B,
                   heads, Nq, float(scale), stream())
    else:
        q2, k2, v2 = (t.reshape(-1, Cc).contiguous() for t in (q, k, v))
        out = torch.full((B * Nq, Cc), float("nan"), dtype=torch.float32, device=DEV)
        _cabi.call("ldmb200_attention_simt", q2.data_ptr(), Cc, 0, k2.data_ptr(), 0, v2.data_ptr(), 0, Cc,
                   out.data_ptr(), Cc, B, heads, Nq, Nkv, float(scale), stream())
    torch.cuda.synchronize()
    return out.float().reshape(B, Nq, Cc)


def attention_reference(q, k, v, heads, scale):
    B, Nq, Cc = q.shape
    d = Cc // heads
    sp = lambda t: t.double().reshape(B, t.shape[1], heads, d).permute(0, 2, 1, 3)
    s = torch.einsum("bhid,bhjd->bhij", sp(q), sp(k)) * scale
    o = torch.einsum("bhij,bhjd->bhid", s.softmax(-1), sp(v))
    return o.permute(0, 2, 1, 3).reshape(B, Nq, Cc).float()


def groupnorm(x1, x2, gamma, beta, B, hw, eps, silu, out_dtype, want_raw=False):
    """x1/x2 channels-last fp32 matrices.  Returns (y, raw)."""
    c = x1.shape[1] + (x2.shape[1] if x2 is not None else 0)
    tdt = TORCH_DT[out_dtype]
    y = torch.full((B * hw, c), float("nan"), dtype=tdt, device=DEV)
    raw = torch.full((B * hw, c), float("nan"), dtype=tdt, device=DEV) if want_raw else None
    lib = _cabi.load()
    ws = torch.empty(int(lib.ldmb200_groupnorm_ws_floats(B, hw, c)), dtype=torch.float32, device=DEV)
    _cabi.call("ldmb200_groupnorm", x1.data_ptr(), x1.shape[1], ptr(x2), x2.shape[1] if x2 is not None else 0,
               gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), ptr(raw), out_dtype, B, hw, 32, float(eps), int(silu),
               ws.data_ptr(), stream())
    torch.cuda.synchronize()
    return y, raw


def groupnorm_fused(x1, cs1, x2, cs2, gamma, beta, B, hw, eps, silu, out_dtype, want_raw=False):
    """Like groupnorm(), statistics from the GEMM epilogues' column sums (ldmb200_gemm_desc.colstats)."""
    c = x1.shape[1] + (x2.shape[1] if x2 is not None else 0)
    tdt = TORCH_DT[out_dtype]
    y = torch.full((B * hw, c), float("nan"), dtype=tdt, device=DEV)
    raw = torch.full((B * hw, c), float("nan"), dtype=tdt, device=DEV) if want_raw else None
    lib = _cabi.load()
    ws = torch.empty(int(lib.ldmb200_groupnorm_ws_floats(B, hw, c)), dtype=torch.float32, device=DEV)
    _cabi.call("ldmb200_groupnorm_fused", x1.data_ptr(), x1.shape[1], cs1.data_ptr(), ptr(x2),
               x2.shape[1] if x2 is not None else 0, ptr(cs2), gamma.data_ptr(), beta.data_ptr(), y.data_ptr(),
               ptr(raw), out_dtype, B, hw, 32, float(eps), int(silu), ws.data_ptr(), stream())
    torch.cuda.synchronize()
    return y, raw


def small_linear_batched(problems):
    """problems: list of (x [R,K], w [N,K], bias [N] or None, silu).  One launch; returns the outputs."""
    lib = _cabi.load()
    arr = (LinearJob * len(problems))()
    outs, first = [], 0
    for j, (x, w, b, silu) in zip(arr, problems):
        rows, k = x.shape
        n = w.shape[0]
        o = torch.full((rows, n), float("nan"), device=DEV)
        outs.append(o)
        j.in_, j.w, j.bias, j.out = x.data_ptr(), w.data_ptr(), ptr(b), o.data_ptr()
        j.in_ld, j.out_ld, j.rows, j.k, j.n, j.silu_in, j.first_block = k, n, rows, k, n, int(silu), first
        first += int(lib.ldmb200_small_linear_blocks(rows, n))
    table = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(DEV)
    _cabi.call("ldmb200_small_linear_batched", table.data_ptr(), len(problems), first, stream())
    torch.cuda.synchronize()
    return outs
