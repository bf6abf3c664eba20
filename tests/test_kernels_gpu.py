"""Per-kernel parity through the C-ABI (``-m gpu``): each sm_100a kernel against a plain PyTorch fp32/fp64
reference of the same op on the same seeded inputs, on the shape classes of SURVEY Appendix A.

Tolerances (rel-L2 = ||a-b||/||b||):
  * fp32 SIMT kernels and norms: 2e-6 .. 1e-5 (summation-order floor);
  * tensor-core kernels: operands are bf16-rounded BEFORE the reference is computed, so what remains is fp32
    accumulation order (<= 1e-5) plus the bf16 rounding of the output when it is stored as bf16 (<= 4e-3);
  * DDIM update: bit-exact.
"""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from dsml_thesis_b200 import _cabi  # noqa: E402
from dsml_thesis_b200._cabi import BF16, F16, F32  # noqa: E402

import kernel_harness as K  # noqa: E402
from oracle import ldm_oracle as O  # noqa: E402


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(0)
    yield
    torch.cuda.synchronize()


def rnd(*shape, scale=1.0):
    return torch.randn(*shape, device="cuda") * scale


def bf(x):
    return x.to(torch.bfloat16).float()


def r16(x, dt):
    return K.rnd16(x, dt)


#: rel-L2 bound of storing an fp32 result in a 16-bit type (half an ulp, uniformly distributed)
OUT_TOL = {F32: 1e-5, BF16: 4e-3, F16: 5e-4}


# ------------------------------------------------------------------------------------------ elementwise
@pytest.mark.parametrize("n", [4, 1000, 2 * 3 * 64 * 64, 64 * 3 * 64 * 64 + 3])
@pytest.mark.parametrize("cfg,noise", [(False, False), (True, False), (False, True), (True, True)])
def test_elementwise_ddim_step_bit_exact(n, cfg, noise):
    from dsml_thesis_b200 import ddim_step
    sch = O.ddim_schedule(O.ddpm_buffers()["alphas_cumprod"], 50, 1.0)
    for index in (49, 25, 0):
        x, ec = rnd(n), rnd(n)
        eu = rnd(n) if cfg else None
        nz = rnd(n) if noise else None
        a_t, a_prev, sg, s1 = (float(sch[k][index]) for k in ("a_t", "a_prev", "sigma_t", "sqrt_one_minus_at"))
        xp, p0 = ddim_step(x, ec, eu, nz, a_t, a_prev, sg, s1, 3.0, 0.7)
        e = ec.cpu() if eu is None else eu.cpu() + 3.0 * (ec.cpu() - eu.cpu())
        rx, rp = O.ddim_update(x.cpu().reshape(1, 1, 1, n), e.reshape(1, 1, 1, n), a_t, a_prev, sg, s1,
                               nz.cpu().reshape(1, 1, 1, n) if nz is not None else None, 0.7)
        assert torch.equal(p0.cpu(), rp.reshape(n)), "pred_x0 not bit-exact"
        assert torch.equal(xp.cpu(), rx.reshape(n)), "x_prev not bit-exact"


def test_elementwise_ddim_step_empty_is_an_error():
    with pytest.raises(_cabi.Ldmb200Error):
        _cabi.call("ldmb200_ddim_step", None, None, None, None, None, None, 0, 1., 1., 0., 0., 1., 1., K.stream())


def test_elementwise_timestep_embedding():
    t = torch.tensor([0, 1, 21, 501, 981, 999], device="cuda")
    for dim in (160, 32):
        out = torch.empty(t.numel(), dim, device="cuda")
        _cabi.call("ldmb200_timestep_embedding", t.data_ptr(), out.data_ptr(), t.numel(), dim, 10000.0, K.stream())
        ref = O.timestep_embedding(t.cpu(), dim)
        # arguments reach ~1e3 rad: one fp32 ulp of the argument is 6e-5, torch's own fp32 sin/cos sit within that
        assert (out.cpu() - ref).abs().max() <= 1.5e-4
        assert K.rel_l2(out, ref) <= 2e-5


@pytest.mark.parametrize("rows,k,n,silu", [(2, 160, 640, 0), (64, 640, 640, 1), (3, 1024, 320, 0), (17, 512, 160, 1),
                                           (33, 64, 100, 0)])
def test_elementwise_small_linear(rows, k, n, silu):
    x, w, b = rnd(rows, k), rnd(n, k, scale=k ** -0.5), rnd(n)
    out = torch.empty(rows, n, device="cuda")
    _cabi.call("ldmb200_small_linear", x.data_ptr(), k, w.data_ptr(), b.data_ptr(), out.data_ptr(), n, rows, k, n,
               silu, K.stream())
    ref = F.linear((F.silu(x) if silu else x).double(), w.double(), b.double()).float()
    assert K.rel_l2(out, ref) <= 2e-6


def test_elementwise_small_linear_batched():
    """One launch over a ragged table of layers (the emb / cross-attention pre-pass of the UNet plan)."""
    shapes = [(64, 640, 160, 1), (64, 640, 640, 1), (2, 512, 320, 0), (64, 320, 320, 0), (1, 64, 32, 0),
              (130, 96, 70, 1), (64, 1024, 640, 0)]
    probs = [(rnd(r, k), rnd(n, k, scale=k ** -0.5), rnd(n) if i % 2 == 0 else None, s)
             for i, (r, k, n, s) in enumerate(shapes)]
    outs = K.small_linear_batched(probs)
    for (x, w, b, s), o in zip(probs, outs):
        ref = F.linear((F.silu(x) if s else x).double(), w.double(), b.double() if b is not None else None).float()
        assert K.rel_l2(o, ref) <= 2e-6


@pytest.mark.parametrize("B,Cc,H,cpad,dt", [(2, 3, 8, 32, BF16), (2, 9, 16, 32, BF16), (3, 3, 8, 3, F32), (2, 3, 8, 32, F16)])
def test_elementwise_layout_roundtrip(B, Cc, H, cpad, dt):
    x = rnd(B, Cc, H, H)
    y = torch.full((B * H * H, cpad), float("nan"), dtype=K.TORCH_DT[dt], device="cuda")
    _cabi.call("ldmb200_nchw_to_nhwc", x.data_ptr(), y.data_ptr(), B, Cc, H * H, cpad, dt, K.stream())
    ref = K.to_cl(x)
    assert torch.equal(y[:, :Cc].float(), r16(ref, dt))
    assert (y[:, Cc:] == 0).all()
    if dt == F32:
        back = torch.empty_like(x)
        _cabi.call("ldmb200_nhwc_to_nchw", y.data_ptr(), cpad, back.data_ptr(), B, Cc, H * H, K.stream())
        assert torch.equal(back, x)


@pytest.mark.parametrize("B,c1,c2,H,eps,silu,dt,raw", [
    (2, 160, 0, 64, 1e-5, 1, BF16, False), (2, 320, 160, 32, 1e-5, 1, BF16, True), (1, 640, 640, 16, 1e-5, 1, F32, True),
    (3, 640, 0, 16, 1e-6, 0, F32, False), (2, 32, 0, 8, 1e-6, 0, BF16, False), (2, 64, 32, 4, 1e-5, 1, F32, True),
    (2, 128, 0, 2, 1e-5, 1, F32, False), (1, 960, 0, 32, 1e-5, 1, BF16, False), (2, 320, 160, 32, 1e-5, 1, F16, True),
    (2, 160, 0, 64, 1e-6, 0, F16, False)])
def test_elementwise_groupnorm(B, c1, c2, H, eps, silu, dt, raw):
    hw = H * H
    x1 = rnd(B * hw, c1) * 2 + 0.5
    x2 = rnd(B * hw, c2) - 1 if c2 else None
    c = c1 + c2
    g, b = 1 + 0.1 * rnd(c), 0.1 * rnd(c)
    y, r = K.groupnorm(x1, x2, g, b, B, hw, eps, silu, dt, raw)
    xc = torch.cat([x1, x2], 1) if c2 else x1
    ref = F.group_norm(K.from_cl(xc, B, H, H).double(), 32, g.double(), b.double(), eps)
    ref = K.to_cl(F.silu(ref) if silu else ref)
    assert K.rel_l2(y, ref) <= (OUT_TOL[dt] if dt != F32 else 3e-6)
    if raw:
        assert torch.equal(r.float(), r16(xc, dt))


@pytest.mark.parametrize("rows,c,dt", [(4096, 160, BF16), (1024, 320, F32), (256, 640, BF16), (5, 32, F32), (77, 1280, F32),
                                       (4096, 160, F16), (300, 640, F16), (4099, 160, F16), (301, 320, BF16), (1, 160, F32),
                                       (33, 320, F32)])
def test_elementwise_layernorm(rows, c, dt):
    x, g, b = rnd(rows, c) * 3 + 1, 1 + 0.1 * rnd(c), 0.1 * rnd(c)
    y = torch.empty(rows, c, dtype=K.TORCH_DT[dt], device="cuda")
    _cabi.call("ldmb200_layernorm", x.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), dt, rows, c, 1e-5, K.stream())
    ref = F.layer_norm(x.double(), (c,), g.double(), b.double(), 1e-5)
    assert K.rel_l2(y, ref) <= (OUT_TOL[dt] if dt != F32 else 2e-6)


def test_elementwise_f16_conversions_saturate():
    """fp16 stores clamp to +-65504 instead of producing inf (LayerNorm with a huge gain as the generator)."""
    rows, c = 8, 64
    x = rnd(rows, c)
    g, b = torch.full((c,), 1e6, device="cuda"), torch.zeros(c, device="cuda")
    y = torch.empty(rows, c, dtype=torch.float16, device="cuda")
    _cabi.call("ldmb200_layernorm", x.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), F16, rows, c, 1e-5, K.stream())
    torch.cuda.synchronize()
    assert torch.isfinite(y.float()).all() and y.float().abs().max() == 65504.0


def test_elementwise_resample_layouts():
    B, Cc, H = 2, 32, 8
    x = rnd(B, Cc, H, H)
    up = torch.empty(B * 4 * H * H, Cc, device="cuda")
    _cabi.call("ldmb200_upsample2x", K.to_cl(x).data_ptr(), up.data_ptr(), F32, B, H, H, Cc, K.stream())
    assert torch.equal(up, K.to_cl(F.interpolate(x, scale_factor=2, mode="nearest")))
    s2d = K.space_to_depth(x, torch.float32).reshape(B, 2, 2, H // 2, H // 2, Cc)
    ref = x.permute(0, 2, 3, 1).reshape(B, H // 2, 2, H // 2, 2, Cc).permute(0, 2, 4, 1, 3, 5)
    assert torch.equal(s2d, ref.contiguous())


# ------------------------------------------------------------------------------------------ weight packer
@pytest.mark.parametrize("dt", [BF16, F16])
@pytest.mark.parametrize("n,c,ks,kind", [(160, 160, 3, 0), (3, 160, 3, 0), (160, 3, 3, 0), (320, 480, 1, 0),
                                         (1280, 160, 1, 1), (256, 32, 1, 1)])
def test_pack_weights(n, c, ks, kind, dt):
    w = rnd(n, c, ks, ks)
    bias = rnd(n)
    wp, fb, n_pad, c_pad = K.pack_weight(w, kind=kind, bias=bias, dtype=dt)
    ref = torch.zeros(n_pad, ks * ks, c_pad, device="cuda")
    src = w
    if kind == 1:
        perm = K.geglu_perm(n, 160 if n % 160 == 0 else 32)
        src = w[perm]
        assert torch.equal(fb, bias[perm])
    ref[:n, :, :c] = src.permute(0, 2, 3, 1).reshape(n, ks * ks, c)
    assert torch.equal(wp.float(), r16(ref, dt))


def test_pack_weights_fused_rows():
    q, k, v = rnd(64, 64), rnd(64, 64), rnd(64, 64)
    wp, _, n_pad, _ = K.pack_weight(None, parts=[q, k, v])
    assert n_pad == 192
    assert torch.equal(wp[:, 0, :].float(), bf(torch.cat([q, k, v])))


# ------------------------------------------------------------------------------------------ SIMT implicit GEMM
@pytest.mark.parametrize("B,cin,cout,H,ks,stride,up", [
    (2, 3, 160, 16, 3, 1, 0), (1, 160, 160, 16, 3, 1, 0), (2, 64, 64, 8, 3, 2, 0), (2, 64, 32, 4, 3, 1, 1),
    (2, 96, 64, 8, 1, 1, 0), (1, 160, 3, 16, 3, 1, 0), (3, 32, 70, 6, 3, 1, 0)])
def test_simt_conv(B, cin, cout, H, ks, stride, up):
    x, w, b = rnd(B, cin, H, H), rnd(cout, cin, ks, ks, scale=(cin * ks * ks) ** -0.5), rnd(cout)
    out = K.igemm("simt", K.to_cl(x), B, H, H, w, b, ksize=ks, stride=stride, upsample=up)
    ref = K.igemm_reference(x, w, b, stride=stride, upsample=up)
    assert K.rel_l2(K.from_cl(out, B, ref.shape[2], ref.shape[3]), ref) <= 2e-6


def test_simt_epilogues_and_concat():
    B, c1, c2, cout, H = 2, 64, 32, 96, 8
    x1, x2 = rnd(B, c1, H, H), rnd(B, c2, H, H)
    w, b = rnd(cout, c1 + c2, 3, 3, scale=0.05), rnd(cout)
    rv, res = rnd(B, cout), rnd(B, cout, H, H)
    out = K.igemm("simt", K.to_cl(x1), B, H, H, w, b, ksize=3, a2=K.to_cl(x2), rowvec=rv, residual=K.to_cl(res))
    ref = K.igemm_reference(torch.cat([x1, x2], 1), w, b, rowvec=rv, residual_nchw=res)
    assert K.rel_l2(K.from_cl(out, B, H, H), ref) <= 2e-6
    # GEGLU linear (attention.py:43-44) and the NCHW store of UNetModel.out
    wg, bg = rnd(256, c1, 1, 1, scale=0.1), rnd(256)
    out = K.igemm("simt", K.to_cl(x1), B, H * H, 1, wg, bg, geglu=True)
    ref = K.igemm_reference(x1, wg, bg, geglu=True)
    assert K.rel_l2(K.from_cl(out, B, H, H), ref) <= 3e-6
    wo, bo = rnd(3, c1, 3, 3, scale=0.05), rnd(3)
    out = K.igemm("simt", K.to_cl(x1), B, H, H, wo, bo, ksize=3, out_nchw=True)
    assert K.rel_l2(out, K.igemm_reference(x1, wo, bo)) <= 2e-6


# ------------------------------------------------------------------------------------------ tcgen05 implicit GEMM
@pytest.mark.parametrize("M,k,n", [(128, 32, 32), (256, 64, 32), (128, 64, 160), (4096, 160, 160), (1024, 320, 960),
                                   (256, 2560, 640), (300, 96, 64), (8192, 160, 1280)])
@pytest.mark.parametrize("dt", [BF16, F16])
def test_tc_gemm_plain(M, k, n, dt):
    a = rnd(M, k).to(K.TORCH_DT[dt])
    w, b = rnd(n, k, 1, 1, scale=k ** -0.5), rnd(n)
    out = K.igemm("tc", a, 1, M, 1, w, b)
    ref = F.linear(a.double(), r16(w.reshape(n, k), dt).double(), b.double()).float()
    assert K.rel_l2(out, ref) <= 1e-5


@pytest.mark.parametrize("M,k,n,odt", [(4096, 160, 160, F32), (4096, 160, 160, F16), (1000, 320, 320, BF16),
                                       (8192, 640, 160, F32), (260, 64, 480, F16)])
def test_tc_gemm_residual_rowvec_every_out_dtype(M, k, n, odt):
    """The coalesced epilogue: bias + per-sample row vector + fp32 residual, fp32 / bf16 / f16 stores, M not a
    multiple of the tile, samples of 4 / 100 / 4096 rows (uniform and per-row row-vector lookup)."""
    for rows_per_sample in (4096, 100, 4):
        if M % rows_per_sample:
            continue
        Bn = M // rows_per_sample
        a = rnd(M, k).to(torch.float16)
        w, b = rnd(n, k, 1, 1, scale=k ** -0.5), rnd(n)
        rv, res = rnd(Bn, n), rnd(M, n)
        out = K.igemm("tc", a, Bn, rows_per_sample, 1, w, b, rowvec=rv, residual=res, out_dtype=odt)
        ref = F.linear(a.double(), r16(w.reshape(n, k), F16).double(), b.double()).float()
        ref = ref + rv.repeat_interleave(rows_per_sample, 0) + res
        assert K.rel_l2(out, ref) <= OUT_TOL[odt], rows_per_sample


@pytest.mark.parametrize("B,cin,cout,H", [(2, 32, 32, 16), (1, 160, 160, 64), (2, 320, 320, 32), (4, 64, 64, 8),
                                          (3, 64, 32, 4), (5, 128, 64, 2), (1, 480, 160, 16), (1, 32, 16, 128)])
def test_tc_conv3x3(B, cin, cout, H):
    x = bf(rnd(B, cin, H, H))
    w, b = rnd(cout, cin, 3, 3, scale=(9 * cin) ** -0.5), rnd(cout)
    out = K.igemm("tc", K.to_cl(x, torch.bfloat16), B, H, H, w, b, ksize=3)
    ref = K.igemm_reference(x, bf(w), b)
    assert K.rel_l2(K.from_cl(out, B, H, H), ref) <= 1e-5


@pytest.mark.parametrize("B,c,H", [(2, 64, 16), (1, 160, 64), (3, 320, 8)])
def test_tc_conv3x3_stride2(B, c, H):
    x = bf(rnd(B, c, H, H))
    w, b = rnd(c, c, 3, 3, scale=(9 * c) ** -0.5), rnd(c)
    out = K.igemm("tc", K.space_to_depth(x, torch.bfloat16), B, H, H, w, b, ksize=3, stride=2)
    ref = K.igemm_reference(x, bf(w), b, stride=2)
    assert K.rel_l2(K.from_cl(out, B, H // 2, H // 2), ref) <= 1e-5


def test_tc_epilogues():
    B, c, H, n = 2, 64, 16, 96
    x = bf(rnd(B, c, H, H))
    a = K.to_cl(x, torch.bfloat16)
    w, b = rnd(n, c, 3, 3, scale=0.05), rnd(n)
    rv, res = rnd(B, n), rnd(B, n, H, H)
    out = K.igemm("tc", a, B, H, H, w, b, ksize=3, rowvec=rv, residual=K.to_cl(res))
    ref = K.igemm_reference(x, bf(w), b, rowvec=rv, residual_nchw=res)
    assert K.rel_l2(K.from_cl(out, B, H, H), ref) <= 1e-5
    for odt in (BF16, F16):
        out = K.igemm("tc", a, B, H, H, w, b, ksize=3, rowvec=rv, residual=K.to_cl(res), out_dtype=odt)
        assert K.rel_l2(K.from_cl(out, B, H, H), ref) <= OUT_TOL[odt]
    # GEGLU with both N tiles (160: C = 160 -> n = 1280; 32: C = 32 -> n = 256)
    for cc in (160, 32):
        xg = bf(rnd(512, cc))
        wg, bg = rnd(8 * cc, cc, 1, 1, scale=cc ** -0.5), rnd(8 * cc)
        y = F.linear(xg.double(), bf(wg.reshape(8 * cc, cc)).double(), bg.double())
        va, ga = y.chunk(2, dim=1)
        for odt in (BF16, F16, F32):
            out = K.igemm("tc", xg.to(torch.bfloat16), 1, 512, 1, wg, bg, geglu=True, out_dtype=odt)
            assert K.rel_l2(out, (va * F.gelu(ga)).float()) <= OUT_TOL[odt]
    # UNetModel.out: n = 3 padded to 16, fp32 NCHW store
    wo, bo = rnd(3, c, 3, 3, scale=0.05), rnd(3)
    out = K.igemm("tc", a, B, H, H, wo, bo, ksize=3, out_nchw=True)
    assert K.rel_l2(out, K.igemm_reference(x, bf(wo), bo)) <= 1e-5


@pytest.mark.parametrize("B,H,c1,c2,n,ksize,stride,extras", [
    (40, 32, 160, 0, 160, 3, 1, "res+rowvec+stats"),    # M = 40960: 160 pairs on 74 clusters, K = 1440 with a partial last chunk
    (40, 32, 320, 0, 320, 3, 1, "res+stats"),           # two N tiles per pair, K = 2880
    (20, 32, 320, 160, 320, 3, 1, "rowvec+stats"),      # virtual concat (second tensor map), K = 4320
    (40, 64, 160, 0, 160, 3, 2, "stats"),               # stride 2 on the space-to-depth planes
    (40, 16, 2560, 0, 640, 1, 1, "res+f16out"),         # ff.net.2 of the 640 level: linear, K = 2560, 16-bit output
    (1, 101, 1280, 0, 640, 1, 1, "res"),                # M = 10201: ragged last tile of the last pair
    (40, 16, 640, 0, 640, 1, 1, "res+stats"),           # proj_out of the 640 level: K = 640, the shallowest reduction taken
])
def test_tc_cta_pair(B, H, c1, c2, n, ksize, stride, extras):
    """The CTA-pair main loop (cta_group::2, M = 256 per MMA, half of the weight tile per CTA; taken for K >= 640 on an even
    number of M tiles filling at least two waves of the SMs) against the float64 convolution of the same bf16-rounded operands, every fused epilogue term on."""
    x1, x2 = bf(rnd(B, c1, H, H)), (bf(rnd(B, c2, H, H)) if c2 else None)
    c = c1 + c2
    w, b = rnd(n, c, ksize, ksize, scale=(ksize * ksize * c) ** -0.5), rnd(n)
    Ho = H // stride
    M = B * Ho * Ho
    tiles_m = (M + 127) // 128
    assert tiles_m % 2 == 0 and ksize * ksize * c >= 640 and tiles_m * (n // 160) >= 2 * 148, "shape would not take the CTA-pair path"
    kw = dict(ksize=ksize, stride=stride)
    res = rnd(B, n, Ho, Ho) + 0.3 if "res" in extras else None
    rv = rnd(B, n) if "rowvec" in extras else None
    cs = torch.full((M // 32, n, 2), float("nan"), device=K.DEV) if "stats" in extras else None
    odt = F16 if "f16out" in extras else F32
    if stride == 2:
        a = K.space_to_depth(x1, torch.bfloat16)
    else:
        a = K.to_cl(x1, torch.bfloat16)
    a2 = K.to_cl(x2, torch.bfloat16) if c2 else None
    outs = [K.igemm("tc", a, B, H, H, w, b, a2=a2, rowvec=rv, residual=K.to_cl(res) if res is not None else None,
                    colstats=cs, out_dtype=odt, **kw) for _ in range(2)]
    assert torch.equal(outs[0], outs[1])
    xin = torch.cat([x1, x2], 1) if c2 else x1
    ref = K.to_cl(K.igemm_reference(xin, bf(w), b, stride=stride, rowvec=rv, residual_nchw=res))
    assert K.rel_l2(outs[0].float(), ref) <= (OUT_TOL[F16] if odt == F16 else 1e-5)
    if cs is not None:
        blk = outs[0].reshape(M // 32, 32, n).double()
        assert K.rel_l2(cs[..., 0], blk.sum(1).float()) <= 1e-5
        assert K.rel_l2(cs[..., 1], (blk * blk).sum(1).float()) <= 1e-5


@pytest.mark.parametrize("rows,cc", [(512, 160), (200, 160), (20480, 320), (64, 640), (10240, 640), (10200, 320)])
def test_tc_geglu_tma_store(rows, cc):
    """GEGLU epilogue with a 16-bit output (TMA store of each warp's 32 x 16 chunk, 32B-swizzled staging) == the fp32-output
    epilogue (per-lane stores) rounded to the same type, bit for bit; ragged last tile clipped (harness guard rows)."""
    xg = bf(rnd(rows, cc))
    wg, bg = rnd(8 * cc, cc, 1, 1, scale=cc ** -0.5), rnd(8 * cc)
    out32 = K.igemm("tc", xg.to(torch.bfloat16), 1, rows, 1, wg, bg, geglu=True, out_dtype=F32)
    y = F.linear(xg.double(), bf(wg.reshape(8 * cc, cc)).double(), bg.double())
    va, ga = y.chunk(2, dim=1)
    assert K.rel_l2(out32, (va * F.gelu(ga)).float()) <= OUT_TOL[F32]
    for odt, tdt in ((BF16, torch.bfloat16), (F16, torch.float16)):
        out = K.igemm("tc", xg.to(torch.bfloat16), 1, rows, 1, wg, bg, geglu=True, out_dtype=odt)
        assert torch.equal(out, out32.to(tdt))


@pytest.mark.parametrize("B,H,c,n,ksize,bias,rowvec,res", [
    (2, 16, 160, 480, 1, True, False, False),      # fused QKV of the 160-channel level
    (3, 8, 160, 480, 1, True, False, False),       # M = 192: ragged last tile, rows 192..255 must be clipped by the TMA store
    (1, 8, 320, 960, 1, False, False, False),      # M = 64 < one tile, no bias
    (2, 16, 96, 160, 3, True, True, False),        # 3x3 convolution + per-sample row vector (hw = 256: uniform per tile)
    (3, 8, 96, 160, 3, True, True, False),         # hw = 64: a tile spans two samples -> the launcher keeps the per-lane store path
    (5, 16, 640, 1920, 1, True, False, False),     # M = 1280, 12 N tiles: several tiles per CTA through both staging tiles
    (2, 16, 160, 160, 1, True, True, True),        # attn1.to_out: + cross-attention row vector + fp32 residual (TMA-loaded ring)
    (3, 8, 640, 160, 1, True, False, True),        # ff.net.2 of the 160 level, ragged last tile: residual rows past M are zero-filled
    (40, 32, 320, 320, 1, True, True, True),       # M = 40960 = 320 x 2 tiles on 148 CTAs: the ring wraps across tiles
    (1, 8, 2560, 640, 1, True, False, True),       # ff.net.2 of the 640 level: K > 640 stays on the per-lane path
])
def test_tc_tma_store_epilogue(B, H, c, n, ksize, bias, rowvec, res):
    """Outputs with only a bias / per-tile row vector (and, for linears, an fp32 residual) to add and no statistics take the
    TMA-store epilogue (igemm_tc_kernel<..., TS[, RS]>): fp32 bit-identical to the per-lane store path (forced by asking for
    column statistics), 16-bit bit-identical to that rounded to the same type, nothing written past row M (guard rows)."""
    hw = H * H
    a = bf(rnd(B * hw, c)).to(torch.bfloat16)
    w, b = rnd(n, c, ksize, ksize, scale=(ksize * ksize * c) ** -0.5), (rnd(n) if bias else None)
    rv = rnd(B, n) if rowvec else None
    rs = rnd(B * hw, n) + 0.25 if res else None
    cs = torch.empty(B * hw // 32, n, 2, device=a.device)
    old32 = K.igemm("tc", a, B, H, H, w, b, ksize=ksize, rowvec=rv, residual=rs, colstats=cs)      # per-lane store path
    new32 = K.igemm("tc", a, B, H, H, w, b, ksize=ksize, rowvec=rv, residual=rs)
    assert torch.equal(new32, old32)
    tgt = K.to_cl(K.igemm_reference(K.from_cl(a.float(), B, H, H), bf(w), b, rowvec=rv,
                                    residual_nchw=K.from_cl(rs, B, H, H) if res else None))
    assert K.rel_l2(new32, tgt) <= 1e-5
    for odt, tdt in ((BF16, torch.bfloat16), (F16, torch.float16)):
        out = K.igemm("tc", a, B, H, H, w, b, ksize=ksize, rowvec=rv, residual=rs, out_dtype=odt)
        assert out.dtype == tdt and torch.equal(out, old32.to(tdt))


@pytest.mark.parametrize("B,H,c,n,ksize,extras", [
    (1, 16, 640, 640, 3, "res+rowvec+stats"),    # M = 256, K = 5760: the FR UNet's 8x8..16x16 levels at batch 1
    (1, 32, 320, 320, 3, "res"),                 # M = 1024, K = 2880
    (4, 16, 288, 160, 3, "stats"),               # M = 1024, K = 2592: partial last channel chunk (288 = 4 x 64 + 32) inside a split
    (1, 16, 2560, 640, 1, "res+f16out"),         # ff.net.2 at batch 1: K = 2560, 16-bit output
    (3, 8, 1280, 1280, 3, "rowvec"),             # M = 192 (ragged last tile), K = 11520, 2 x 8 tiles
    (1, 16, 2880, 640, 1, "plain"),              # plain 1x1, K = 2880
])
def test_tc_split_k(B, H, c, n, ksize, extras):
    """ldmb200_gemm_desc.splitk_ws: the split-K launch (small grids, deep K) equals the single-CTA-per-tile launch up to
    fp32 summation order, is bit-reproducible run to run, leaves its counters zero, and never writes outside the
    workspace it was given."""
    hw = H * H
    M = B * hw
    a = bf(rnd(M, c)).to(torch.bfloat16)
    w, b = rnd(n, c, ksize, ksize, scale=(ksize * ksize * c) ** -0.5), rnd(n)
    kw = dict(ksize=ksize)
    if "res" in extras:
        kw["residual"] = rnd(M, n) + 0.3
    if "rowvec" in extras:
        kw["rowvec"] = rnd(B, n)
    if "f16out" in extras:
        kw["out_dtype"] = F16
    mk_cs = lambda: torch.full((M // 32, n, 2), float("nan"), device=a.device) if "stats" in extras else None
    cs0 = mk_cs()
    ref = K.igemm("tc", a, B, H, H, w, b, colstats=cs0, **kw)
    words = (1024 + (16 << 20)) // 4
    guard = torch.full((words + 1024,), 777.0, device=a.device)
    ws = guard[:words]
    ws.zero_()
    outs = []
    for _ in range(3):
        cs = mk_cs()
        outs.append((K.igemm("tc", a, B, H, H, w, b, colstats=cs, splitk_ws=ws, **kw), cs))
        assert bool((ws[:256].view(torch.int32) == 0).all()), "arrival counters must be left at zero"
    assert bool((guard[words:] == 777.0).all())
    assert float(ws[256:].abs().max()) > 0.0, "the shape was expected to take the split-K path (partials untouched)"
    out, cs = outs[0]
    assert torch.isfinite(out.float()).all()
    assert K.rel_l2(out, ref) <= (2e-3 if "f16out" in extras else 1e-5)
    for o2, c2 in outs[1:]:
        assert torch.equal(o2, out)
        if cs is not None:
            assert torch.equal(c2, cs)
    if cs is not None:
        blk = out.reshape(M // 32, 32, n).double()
        assert K.rel_l2(cs[..., 0], blk.sum(1).float()) <= 1e-5
        assert K.rel_l2(cs[..., 1], (blk * blk).sum(1).float()) <= 1e-5
    x = K.from_cl(a.float(), B, H, H)
    tgt = K.igemm_reference(x, bf(w), b, rowvec=kw.get("rowvec"),
                            residual_nchw=K.from_cl(kw["residual"], B, H, H) if "residual" in kw else None)
    assert K.rel_l2(out.float(), K.to_cl(tgt).float()) <= (2e-3 if "f16out" in extras else 1e-5)


def test_tc_split_k_not_taken_without_workspace_or_on_large_grids():
    """No workspace -> never split; a grid that already fills the SMs ignores the workspace (partials stay untouched)."""
    B, H, c, n = 8, 32, 320, 320     # M = 8192 -> 64 x 2 tiles
    a = bf(rnd(B * H * H, c)).to(torch.bfloat16)
    w, b = rnd(n, c, 3, 3, scale=(9 * c) ** -0.5), rnd(n)
    ws = torch.zeros((1024 + (16 << 20)) // 4, device=a.device)
    ref = K.igemm("tc", a, B, H, H, w, b, ksize=3)
    out = K.igemm("tc", a, B, H, H, w, b, ksize=3, splitk_ws=ws)
    assert torch.equal(out, ref)
    assert float(ws.abs().max()) == 0.0


@pytest.mark.parametrize("B,H,c1,c2,n1,n2", [(2, 16, 64, 0, 160, 0), (2, 8, 96, 64, 640, 320), (1, 32, 32, 48, 320, 160),
                                                  (3, 8, 32, 0, 64, 0)])
def test_tc_colstats_groupnorm_fused(B, H, c1, c2, n1, n2):
    """GEMM epilogue column statistics (ldmb200_gemm_desc.colstats) -> ldmb200_groupnorm_fused == GroupNorm of the
    stored fp32 tensors (incl. a virtual concat whose group boundaries straddle the two tensors: 640+320 / 32)."""
    hw = H * H
    outs, css = [], []
    for c, n in ((c1, n1), (c2, n2)):
        if not n:
            continue
        a = bf(rnd(B * hw, c)).to(torch.bfloat16)
        w, b = rnd(n, c, 3, 3, scale=(9 * c) ** -0.5), rnd(n)
        res = rnd(B * hw, n) + 0.7
        guard = torch.full(((B * hw // 32 + 4) * n * 2,), 12345.0, device=a.device)   # M = 192: no write past the end
        cs = guard[: B * hw // 32 * n * 2].view(B * hw // 32, n, 2)
        cs.fill_(float("nan"))
        out = K.igemm("tc", a, B, H, H, w, b, ksize=3, residual=res, colstats=cs)
        assert bool((guard[B * hw // 32 * n * 2:] == 12345.0).all())
        blk = out.reshape(B * hw // 32, 32, n).double()
        assert K.rel_l2(cs[..., 0], blk.sum(1).float()) <= 1e-5
        assert K.rel_l2(cs[..., 1], (blk * blk).sum(1).float()) <= 1e-5
        outs.append(out)
        css.append(cs)
    x1, x2 = outs[0], (outs[1] if len(outs) > 1 else None)
    c = n1 + n2
    g, be = 1 + 0.1 * rnd(c), 0.1 * rnd(c)
    for dt in (F16, BF16):
        y, r = K.groupnorm_fused(x1, css[0], x2, css[1] if x2 is not None else None, g, be, B, hw, 1e-5, 1, dt, True)
        xc = torch.cat([x1, x2], 1) if x2 is not None else x1
        ref = K.to_cl(F.silu(F.group_norm(K.from_cl(xc, B, H, H).double(), 32, g.double(), be.double(), 1e-5)))
        assert K.rel_l2(y, ref) <= OUT_TOL[dt]
        assert torch.equal(r.float(), r16(xc, dt))


# ------------------------------------------------------------------------------------------ attention
@pytest.mark.parametrize("B,N,M,heads", [(2, 64, 64, 2), (1, 256, 256, 5), (2, 100, 7, 3), (1, 130, 1, 1), (2, 16, 77, 4)])
def test_attn_simt(B, N, M, heads):
    c = heads * 32
    q, k, v = rnd(B, N, c), rnd(B, M, c), rnd(B, M, c)
    out = K.attention("simt", q, k, v, heads, 32 ** -0.5)
    assert K.rel_l2(out, K.attention_reference(q, k, v, heads, 32 ** -0.5)) <= 3e-6


@pytest.mark.parametrize("B,N,heads,scale_in", [(1, 128, 1, 1.0), (2, 256, 2, 1.0), (1, 1024, 10, 1.0), (2, 64, 4, 1.0),
                                                (3, 16, 1, 1.0), (1, 4, 20, 1.0), (1, 4096, 5, 1.0), (2, 320, 2, 3.0),
                                                (2, 512, 3, 1.0), (3, 200, 2, 2.0), (1, 129, 1, 1.0)])
@pytest.mark.parametrize("odt", [BF16, F16])
def test_attn_tc(B, N, heads, scale_in, odt):
    c = heads * 32
    q, k, v = (bf(rnd(B, N, c) * scale_in) for _ in range(3))
    out = K.attention("tc", q, k, v, heads, 32 ** -0.5, out_dtype=odt)
    ref = K.attention_reference(q, k, v, heads, 32 ** -0.5)
    # P is rounded to bf16 before PV; the output is stored as bf16 / f16
    assert K.rel_l2(out, ref) <= (6e-3 if odt == BF16 else 4e-3)


@pytest.mark.parametrize("N", [192, 1024, 1100])
@pytest.mark.parametrize("growth", [40.0, 400.0])
def test_attn_tc_rereference(N, growth):
    """Scores that keep growing along the keys: the row maximum of every later 64-key step exceeds the row's
    running reference by far more than 2^8, so the kernel must re-reference (rescale the TMEM-resident output in
    place and recompute the step's P) many times per row — the overflow-safety path of the single-pass softmax."""
    B, heads, c = 2, 2, 64
    q = bf(rnd(B, N, c))
    ramp = torch.linspace(0.0, 1.0, N, device=q.device)[None, :, None]
    k = bf(rnd(B, N, c) + growth * ramp * torch.sign(rnd(1, 1, c)))
    q = bf(q.abs() * torch.sign(k[:, -1:, :]))     # every query is aligned with the ramp direction
    v = bf(rnd(B, N, c))
    out = K.attention("tc", q, k, v, heads, 32 ** -0.5, out_dtype=F16)
    ref = K.attention_reference(q, k, v, heads, 32 ** -0.5)
    assert torch.isfinite(out).all()
    assert K.rel_l2(out, ref) <= 6e-3


# ------------------------------------------------------------------------------------------ K10 folded into the conv
@pytest.mark.parametrize("B,H,cin,cout", [(2, 16, 32, 32), (1, 32, 64, 160), (3, 8, 160, 320), (2, 4, 32, 64)])
@pytest.mark.parametrize("dt", [BF16, F16])
def test_tc_upsample_fold(B, H, cin, cout, dt):
    """Upsample.forward (openaimodel.py:109-119) on the tensor-core path: nearest-2x + 3x3 conv as four 2x2 convolutions on
    the low-res map with tap-summed weights (pack kind 4), against F.interpolate + conv2d on the same 16-bit-rounded
    input; the epilogue's GroupNorm column statistics must describe the tensor it wrote."""
    x = r16(rnd(B, cin, H, H), dt)
    w, b = rnd(cout, cin, 3, 3) / (3 * cin ** 0.5), 0.1 * rnd(cout)
    a = K.to_cl(x).to(K.TORCH_DT[dt])
    M = B * 4 * H * H
    cs = torch.full((M // 32, cout, 2), float("nan"), device="cuda") if (H * H) % 32 == 0 else None
    out = K.igemm("tc", a, B, H, H, w, b, ksize=3, upsample=1, colstats=cs)
    ref = K.to_cl(K.igemm_reference(x, w, b, upsample=1))
    # the four tap sums are rounded to 16 bits once each (instead of nine separately rounded taps): operand-rounding noise
    assert K.rel_l2(out, ref) <= OUT_TOL[dt]
    if cs is not None:
        o = out.reshape(B, 4 * H * H, cout).double()
        tot = cs.reshape(B, 4 * H * H // 32, cout, 2).double().sum(1)
        assert torch.allclose(tot[..., 0], o.sum(1), rtol=1e-4, atol=1e-3)
        assert torch.allclose(tot[..., 1], (o * o).sum(1), rtol=1e-4, atol=1e-3)
