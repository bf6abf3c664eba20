"""GPU parity of the training-step kernels through the C-ABI: every backward kernel against torch
autograd in float64 on the same (bf16-rounded where the kernel takes bf16) inputs."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from dsml_thesis_b200 import _cabi
from dsml_thesis_b200._cabi import BF16, F32, GemmDesc, PackJob, WgradDesc
import kernel_harness as K  # noqa: E402
from kernel_harness import DEV, ptr, rel_l2, stream, to_cl  # noqa: E402

pytestmark = pytest.mark.gpu


def g(seed=0):
    return torch.Generator(device=DEV).manual_seed(seed)


def randn(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, device=DEV, generator=g(seed)) * scale


def bf(x):
    return x.to(torch.bfloat16)


# ---------------------------------------------------------------------------------------------
def test_q_sample_bit_exact_and_mse():
    B, n = 5, 3 * 16 * 16
    x0, nz = randn(B, n, seed=1), randn(B, n, seed=2)
    t = torch.tensor([0, 999, 17, 500, 3], device=DEV)
    sa, s1 = torch.rand(1000, device=DEV, generator=g(3)), torch.rand(1000, device=DEV, generator=g(4))
    out = torch.empty_like(x0)
    _cabi.call("ldmb200_q_sample", x0.data_ptr(), nz.data_ptr(), t.data_ptr(), sa.data_ptr(), s1.data_ptr(),
               out.data_ptr(), B, n, stream())
    ref = sa[t][:, None] * x0 + s1[t][:, None] * nz
    assert torch.equal(out, ref)
    pred = randn(B, n, seed=5)
    loss, dpred = torch.empty(B, device=DEV), torch.empty_like(pred)
    _cabi.call("ldmb200_mse_loss_grad", pred.data_ptr(), nz.data_ptr(), loss.data_ptr(), dpred.data_ptr(), B, n, stream())
    p64 = pred.double().requires_grad_()
    l64 = ((p64 - nz.double()) ** 2).mean(1)
    l64.mean().backward()
    assert rel_l2(loss, l64) < 1e-6
    assert rel_l2(dpred, p64.grad) < 1e-6


@pytest.mark.parametrize("in_dt", [F32, BF16])
@pytest.mark.parametrize("B,hw,n,nv", [(2, 64, 160, 160), (3, 16, 16, 3), (2, 256, 1280, 1280), (1, 40, 32, 32)])
def test_grad_cast_and_column_sums(B, hw, n, nv, in_dt):
    x = randn(B * hw, n, seed=7)
    src = bf(x) if in_dt == BF16 else x
    g16 = torch.full((B * hw, n), float("nan"), dtype=torch.bfloat16, device=DEV)
    ps = torch.zeros(B, n + 4, device=DEV)
    tot, tot2 = torch.ones(nv, device=DEV), torch.zeros(nv, device=DEV)
    _cabi.call("ldmb200_grad_cast", src.data_ptr(), in_dt, g16.data_ptr(), B * hw, n, hw, ps.data_ptr(), n + 4,
               tot.data_ptr(), tot2.data_ptr(), nv, stream())
    ref = src.double()
    assert torch.equal(g16, src.to(torch.bfloat16))
    assert rel_l2(ps[:, :nv], ref.reshape(B, hw, n).sum(1)[:, :nv]) < 1e-5
    assert float(ps[:, nv:].abs().max()) == 0.0
    assert rel_l2(tot - 1.0, ref.sum(0)[:nv]) < 1e-4
    assert rel_l2(tot2, ref.sum(0)[:nv]) < 1e-5


# ---------------------------------------------------------------------------------------------
def run_wgrad(dy16, x16, n, c, B, h, w, ks, stride, c_ld=None):
    taps = ks * ks
    dw = torch.zeros(n, c, ks, ks, device=DEV)
    d = WgradDesc()
    d.dy, d.x, d.dw = dy16.data_ptr(), x16.data_ptr(), dw.data_ptr()
    d.n_stride, d.c_stride, d.tap_stride = c * taps, taps, 1
    d.dy_ld, d.n, d.c_ld, d.c = dy16.shape[1], n, (c_ld or x16.shape[1]), c
    d.batch, d.h, d.w, d.ksize, d.stride = B, h, w, ks, stride
    _cabi.call("ldmb200_wgrad", C.byref(d), stream())
    torch.cuda.synchronize()
    return dw


@pytest.mark.parametrize("B,cin,cout,H,ks,stride", [
    (2, 32, 32, 8, 1, 1), (1, 160, 160, 16, 1, 1), (2, 64, 96, 8, 3, 1), (1, 160, 320, 16, 3, 1),
    (3, 32, 64, 4, 3, 1), (2, 64, 64, 16, 3, 2), (1, 320, 160, 8, 3, 1), (2, 640, 160, 4, 1, 1)])
def test_wgrad_vs_autograd(B, cin, cout, H, ks, stride):
    x = bf(randn(B, cin, H, H, seed=11)).float()
    Ho = H // stride
    dy = bf(randn(B, cout, Ho, Ho, seed=12)).float()
    w = randn(cout, cin, ks, ks, seed=13).double().requires_grad_()
    y = F.conv2d(x.double(), w, None, stride=stride, padding=ks // 2)
    y.backward(dy.double())
    x16 = K.space_to_depth(x, torch.bfloat16) if stride == 2 else bf(to_cl(x))
    dw = run_wgrad(bf(to_cl(dy)), x16, cout, cin, B, Ho, Ho, ks, stride)
    assert rel_l2(dw, w.grad) < 2e-5, rel_l2(dw, w.grad)


def test_wgrad_padded_operands_and_accumulation():
    # the 3-channel input conv (x padded to 32 channels) and the 3-channel output conv (dy padded to 16 columns)
    B, H = 2, 8
    x = bf(randn(B, 3, H, H, seed=21)).float()
    dy = bf(randn(B, 160, H, H, seed=22)).float()
    w = randn(160, 3, 3, 3, seed=23).double().requires_grad_()
    F.conv2d(x.double(), w, None, padding=1).backward(dy.double())
    xp = torch.zeros(B * H * H, 32, device=DEV, dtype=torch.bfloat16)
    xp[:, :3] = bf(to_cl(x))
    dw = run_wgrad(bf(to_cl(dy)), xp, 160, 3, B, H, H, 3, 1)
    assert rel_l2(dw, w.grad) < 2e-5
    x2 = bf(randn(B, 160, H, H, seed=24)).float()
    dy2 = bf(randn(B, 3, H, H, seed=25)).float()
    w2 = randn(3, 160, 3, 3, seed=26).double().requires_grad_()
    F.conv2d(x2.double(), w2, None, padding=1).backward(dy2.double())
    dyp = torch.zeros(B * H * H, 16, device=DEV, dtype=torch.bfloat16)
    dyp[:, :3] = bf(to_cl(dy2))
    dw2 = run_wgrad(dyp, bf(to_cl(x2)), 3, 160, B, H, H, 3, 1)
    assert rel_l2(dw2, w2.grad) < 2e-5


def pack_dgrad(w, n_rows_pad=None):
    """fp32 [n, c, k, k] -> kind-3 packed bf16 [c_pad_rows, k*k, n_pad_cols]."""
    n, c, ks = w.shape[0], w.shape[1], w.shape[2]
    rows = (c + 31) // 32 * 32
    cols = 16 if n <= 16 else (n + 31) // 32 * 32
    dst = torch.full((rows, ks * ks, cols), float("nan"), dtype=torch.bfloat16, device=DEV)
    arr = (PackJob * 1)()
    j = arr[0]
    wc = w.contiguous()
    j.src, j.dst, j.kind, j.n, j.c, j.ksize = wc.data_ptr(), dst.data_ptr(), 3, n, c, ks
    j.n_pad, j.c_pad, j.row_offset, j.block_n, j.dst_dtype = rows, cols, 0, cols, BF16
    dev = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(DEV)
    _cabi.call("ldmb200_pack_weights", dev.data_ptr(), 1, stream())
    torch.cuda.synchronize()
    return dst, rows, cols


def run_dgrad(dy16, wt, rows, B, h, w, ks, residual=None, out_dtype=F32):
    d = GemmDesc()
    d.a, d.c1, d.a_dtype = dy16.data_ptr(), dy16.shape[1], BF16
    d.batch, d.h_in, d.w_in, d.ksize, d.stride = B, h, w, ks, 1
    d.w, d.n = wt.data_ptr(), rows
    d.residual, d.residual_ld = ptr(residual), (residual.shape[1] if residual is not None else 0)
    out = torch.full((B * h * w, rows), float("nan"), dtype=K.TORCH_DT[out_dtype], device=DEV)
    d.out, d.out_dtype, d.out_ld, d.n_valid = out.data_ptr(), out_dtype, rows, rows
    _cabi.call("ldmb200_igemm_tc", C.byref(d), stream())
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("B,cin,cout,H,ks", [(2, 160, 160, 8, 3), (1, 320, 160, 16, 3), (2, 160, 320, 8, 1),
                                             (2, 160, 3, 8, 3), (1, 640, 1920, 16, 1)])
def test_dgrad_through_igemm_with_transposed_pack(B, cin, cout, H, ks):
    w = bf(randn(cout, cin, ks, ks, seed=31, scale=0.1)).float()
    dy = bf(randn(B, cout, H, H, seed=32)).float()
    x = randn(B, cin, H, H, seed=33).double().requires_grad_()
    F.conv2d(x, w.double(), None, padding=ks // 2).backward(dy.double())
    wt, rows, cols = pack_dgrad(w)
    dyp = torch.zeros(B * H * H, cols, dtype=torch.bfloat16, device=DEV)
    dyp[:, :cout] = bf(to_cl(dy))
    res = randn(B * H * H, rows, seed=34)
    dx = run_dgrad(dyp, wt, rows, B, H, H, ks, residual=res)
    ref = to_cl(x.grad.float()) + res[:, :cin]
    assert rel_l2(dx[:, :cin], ref) < 1e-5


def test_dgrad_stride2_via_zero_insert():
    B, c, H = 2, 160, 16
    w = bf(randn(c, c, 3, 3, seed=41, scale=0.1)).float()
    dy = bf(randn(B, c, H // 2, H // 2, seed=42)).float()
    x = randn(B, c, H, H, seed=43).double().requires_grad_()
    F.conv2d(x, w.double(), None, stride=2, padding=1).backward(dy.double())
    dy16 = bf(to_cl(dy))
    z = torch.full((B * H * H, c), float("nan"), dtype=torch.bfloat16, device=DEV)
    _cabi.call("ldmb200_zero_insert2x", dy16.data_ptr(), z.data_ptr(), B, H // 2, H // 2, c, stream())
    wt, rows, _ = pack_dgrad(w)
    dx = run_dgrad(z, wt, rows, B, H, H, 3)
    assert rel_l2(dx, to_cl(x.grad.float())) < 1e-5


def test_downsum2x():
    B, c, H = 2, 64, 8
    du = randn(B, c, 2 * H, 2 * H, seed=51)
    x = randn(B, c, H, H, seed=52).double().requires_grad_()
    F.interpolate(x, scale_factor=2, mode="nearest").backward(du.double())
    src = to_cl(du)
    out = torch.empty(B * H * H, c, device=DEV)
    _cabi.call("ldmb200_downsum2x", src.data_ptr(), out.data_ptr(), B, H, H, c, stream())
    assert rel_l2(out, to_cl(x.grad.float())) < 1e-6


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,c1,c2,H,eps,silu,add,fused_like", [
    (2, 160, 0, 8, 1e-5, 1, None, 0), (2, 64, 96, 4, 1e-5, 1, "bf16", 0), (1, 320, 0, 16, 1e-6, 0, "f32", 0),
    (3, 640, 640, 4, 1e-5, 1, "bf16", 0), (2, 32, 0, 2, 1e-5, 1, "f32", 0), (2, 160, 160, 8, 1e-5, 1, None, 1)])
def test_groupnorm_backward(B, c1, c2, H, eps, silu, add, fused_like):
    hw, Cc = H * H, c1 + c2
    x1 = randn(B * hw, c1, seed=61) * 2 + 0.5
    x2 = randn(B * hw, c2, seed=62) if c2 else None
    gamma, beta = randn(Cc, seed=63), randn(Cc, seed=64)
    dy = bf(randn(B * hw, Cc, seed=65))
    addend = None
    if add == "bf16":
        addend = bf(randn(B * hw, Cc, seed=66))
    elif add == "f32":
        addend = randn(B * hw, Cc, seed=66)
    lib = _cabi.load()
    ws = torch.empty(int(lib.ldmb200_groupnorm_ws_floats(B, hw, Cc)), device=DEV)
    ws2 = torch.empty_like(ws)
    y = torch.empty(B * hw, Cc, dtype=torch.bfloat16, device=DEV)
    _cabi.call("ldmb200_groupnorm", x1.data_ptr(), c1, ptr(x2), c2, gamma.data_ptr(), beta.data_ptr(), y.data_ptr(),
               None, BF16, B, hw, 32, float(eps), silu, ws.data_ptr(), stream())
    xcat = torch.cat([x1, x2], 1) if c2 else x1
    if fused_like:  # what ldmb200_groupnorm_fused leaves in the workspace: [b][g] = {mean, rstd}
        xg = xcat.double().reshape(B, hw, 32, Cc // 32)
        mean = xg.mean((1, 3))
        var = xg.var((1, 3), unbiased=False)
        ws[: B * 64] = torch.stack([mean, 1.0 / torch.sqrt(var + eps)], -1).float().reshape(-1)
    dx1 = randn(B * hw, c1, seed=67)
    dx1_0 = dx1.clone()
    dx2 = torch.full((B * hw, max(c2, 1)), float("nan"), device=DEV)
    dg, db = torch.zeros(Cc, device=DEV), torch.zeros(Cc, device=DEV)
    _cabi.call("ldmb200_groupnorm_bwd", x1.data_ptr(), c1, ptr(x2), c2, gamma.data_ptr(), beta.data_ptr(),
               dy.data_ptr(), ptr(addend), (BF16 if add == "bf16" else F32), dx1.data_ptr(), 1,
               dx2.data_ptr() if c2 else None, 0, dg.data_ptr(), db.data_ptr(), B, hw, 32, float(eps), silu,
               ws.data_ptr(), fused_like, ws2.data_ptr(), stream())
    x64 = xcat.double().reshape(B, hw, Cc).permute(0, 2, 1).requires_grad_()
    g64, b64 = gamma.double().requires_grad_(), beta.double().requires_grad_()
    o = F.group_norm(x64, 32, g64, b64, eps)
    if silu:
        o = F.silu(o)
    o.backward(dy.double().reshape(B, hw, Cc).permute(0, 2, 1))
    ref = x64.grad.permute(0, 2, 1).reshape(B * hw, Cc)
    if addend is not None:
        ref = ref + addend.double()
    assert rel_l2(dx1 - dx1_0, ref[:, :c1]) < 2e-4, rel_l2(dx1 - dx1_0, ref[:, :c1])
    if c2:
        assert rel_l2(dx2, ref[:, c1:]) < 2e-4
    assert rel_l2(dg, g64.grad) < 2e-4
    assert rel_l2(db, b64.grad) < 2e-4


@pytest.mark.parametrize("rows,c,add", [(4096, 160, True), (1000, 320, False), (256, 640, True), (7, 32, True)])
def test_layernorm_backward(rows, c, add):
    x = randn(rows, c, seed=71) * 1.5 + 0.3
    gamma, dy = randn(c, seed=72), bf(randn(rows, c, seed=73))
    addend = randn(rows, c, seed=74) if add else None
    dx = addend.clone() if add else torch.empty(rows, c, device=DEV)
    dg, db = torch.zeros(c, device=DEV), torch.zeros(c, device=DEV)
    _cabi.call("ldmb200_layernorm_bwd", x.data_ptr(), gamma.data_ptr(), dy.data_ptr(), dx.data_ptr() if add else None,
               dx.data_ptr(), dg.data_ptr(), db.data_ptr(), rows, c, 1e-5, stream())
    x64, g64 = x.double().requires_grad_(), gamma.double().requires_grad_()
    b64 = torch.zeros(c, device=DEV, dtype=torch.float64, requires_grad=True)
    F.layer_norm(x64, (c,), g64, b64, 1e-5).backward(dy.double())
    ref = x64.grad + (addend.double() if add else 0)
    assert rel_l2(dx, ref) < 1e-5
    assert rel_l2(dg, g64.grad) < 1e-4
    assert rel_l2(db, b64.grad) < 1e-4


@pytest.mark.parametrize("rows,inner", [(256, 640), (100, 1280), (4096, 128)])
def test_geglu_forward_backward(rows, inner):
    u = bf(randn(rows, 2 * inner, seed=81))
    dout = bf(randn(rows, inner, seed=82))
    out = torch.empty(rows, inner, dtype=torch.bfloat16, device=DEV)
    du = torch.empty(rows, 2 * inner, dtype=torch.bfloat16, device=DEV)
    _cabi.call("ldmb200_geglu_fwd", u.data_ptr(), out.data_ptr(), rows, inner, stream())
    _cabi.call("ldmb200_geglu_bwd", u.data_ptr(), dout.data_ptr(), du.data_ptr(), rows, inner, stream())
    u64 = u.double().requires_grad_()
    a, gt = u64.chunk(2, -1)
    o = a * F.gelu(gt)
    o.backward(dout.double())
    assert rel_l2(out, o) < 4e-3       # bf16 output rounding
    assert rel_l2(du, u64.grad) < 4e-3


@pytest.mark.parametrize("impl", ["tc", "mma"])
@pytest.mark.parametrize("B,N,heads", [(1, 64, 1), (2, 256, 2), (1, 1024, 5), (2, 16, 3), (1, 100, 2), (1, 200, 1),
                                       (3, 4096, 2), (2, 130, 1)])
def test_attention_backward(B, N, heads, impl):
    Cc = heads * 32
    scale = 32 ** -0.5
    qkv = bf(randn(B * N, 3 * Cc, seed=91))
    dout = bf(randn(B * N, Cc, seed=92))
    out = torch.empty(B * N, Cc, dtype=torch.bfloat16, device=DEV)
    lse = torch.full((B, heads, N), float("nan"), device=DEV)
    _cabi.call("ldmb200_attention_tc_lse", qkv.data_ptr(), 3 * Cc, 0, Cc, 2 * Cc, out.data_ptr(), Cc, BF16,
               lse.data_ptr(), B, heads, N, scale, stream())
    dqkv = torch.full((B * N, 3 * Cc), float("nan"), dtype=torch.bfloat16, device=DEV)
    ws = None
    if impl == "tc":
        ws = torch.empty(int(_cabi.load().ldmb200_attention_bwd_ws_floats(B, heads, N)), device=DEV)
    _cabi.call("ldmb200_attention_bwd", qkv.data_ptr(), 3 * Cc, 0, Cc, 2 * Cc, out.data_ptr(), dout.data_ptr(), Cc,
               lse.data_ptr(), dqkv.data_ptr(), ptr(ws), B, heads, N, scale, stream())
    x = qkv.double().reshape(B, N, 3, heads, 32).requires_grad_()
    q, k, v = (x[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    s = torch.einsum("bhid,bhjd->bhij", q, k) * scale
    o = torch.einsum("bhij,bhjd->bhid", s.softmax(-1), v).permute(0, 2, 1, 3).reshape(B * N, Cc)
    o.backward(dout.double())
    lse_ref = torch.logsumexp(s, -1) * 1.4426950408889634
    assert rel_l2(out, o) < 1e-2
    assert float((lse - lse_ref).abs().max()) < 2e-2
    ref = x.grad.reshape(B * N, 3 * Cc)
    for i, name in enumerate("qkv"):
        err = rel_l2(dqkv[:, i * Cc:(i + 1) * Cc], ref[:, i * Cc:(i + 1) * Cc])
        assert err < 1.5e-2, (name, err)


@pytest.mark.parametrize("rows,k,n,silu", [(2, 160, 640, 0), (32, 640, 320, 1), (3, 1024, 160, 0)])
def test_small_linear_backward(rows, k, n, silu):
    x, w, dout = randn(rows, k, seed=101), randn(n, k, seed=102, scale=0.1), randn(rows, n, seed=103)
    dw, dbias = torch.zeros(n, k, device=DEV), torch.zeros(n, device=DEV)
    din = randn(rows, k, seed=104)
    din0 = din.clone()
    _cabi.call("ldmb200_small_linear_bwd", x.data_ptr(), k, w.data_ptr(), dout.data_ptr(), n, dw.data_ptr(),
               dbias.data_ptr(), din.data_ptr(), k, 1, rows, k, n, silu, stream())
    x64, w64 = x.double().requires_grad_(), w.double().requires_grad_()
    b64 = torch.zeros(n, device=DEV, dtype=torch.float64, requires_grad=True)
    F.linear(F.silu(x64) if silu else x64, w64, b64).backward(dout.double())
    assert rel_l2(dw, w64.grad) < 1e-5
    assert rel_l2(dbias, b64.grad) < 1e-5
    assert rel_l2(din - din0, x64.grad) < 1e-5


def test_small_linear_backward_batched_accumulates_shared_input_gradient():
    from dsml_thesis_b200._cabi import LinearBwdJob
    rows, k = 4, 128
    x = randn(rows, k, seed=121)
    probs = [(randn(n, k, seed=122 + i, scale=0.1), randn(rows, n, seed=130 + i)) for i, n in enumerate((64, 160, 32))]
    din = torch.zeros(rows, k, device=DEV)
    dws = [torch.zeros_like(w) for w, _ in probs]
    dbs = [torch.zeros(w.shape[0], device=DEV) for w, _ in probs]
    arr = (LinearBwdJob * len(probs))()
    for j, (w, dout), dw, db in zip(arr, probs, dws, dbs):
        j.in_, j.w, j.dout, j.dw, j.dbias, j.din = (x.data_ptr(), w.data_ptr(), dout.data_ptr(), dw.data_ptr(),
                                                    db.data_ptr(), din.data_ptr())
        j.in_ld, j.dout_ld, j.din_ld, j.rows, j.k, j.n, j.silu_in = k, w.shape[0], k, rows, k, w.shape[0], 1
    table = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(DEV)
    _cabi.call("ldmb200_small_linear_bwd_batched", table.data_ptr(), len(probs), stream())
    x64 = x.double().requires_grad_()
    ws = [w.double().requires_grad_() for w, _ in probs]
    bs = [torch.zeros(w.shape[0], device=DEV, dtype=torch.float64, requires_grad=True) for w, _ in probs]
    sum((F.linear(F.silu(x64), w, b) * dout.double()).sum() for w, b, (_, dout) in zip(ws, bs, probs)).backward()
    assert rel_l2(din, x64.grad) < 1e-5
    for dw, db, w, b in zip(dws, dbs, ws, bs):
        assert rel_l2(dw, w.grad) < 1e-5 and rel_l2(db, b.grad) < 1e-5


def test_adamw_ema_matches_torch():
    n = 10007
    p0, grad = randn(n, seed=111), randn(n, seed=112)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)
    p, m, v, ema = p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV), p0.clone()
    ema_ref = p0.clone()
    for step in range(1, 4):
        ref.grad = grad * step
        opt.step()
        ema_ref.sub_(0.01 * (ema_ref - ref.detach()))
        gstep = grad * step
        _cabi.call("ldmb200_adamw_ema", p.data_ptr(), gstep.data_ptr(), m.data_ptr(), v.data_ptr(), ema.data_ptr(), n,
                   1e-3, 0.9, 0.999, 1e-8, 1e-2, 1 - 0.9 ** step, 1 - 0.999 ** step, 0.01, stream())
    assert rel_l2(p, ref.detach()) < 1e-6
    assert rel_l2(ema, ema_ref) < 1e-6
