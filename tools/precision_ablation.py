"""CPU ablation: which 16-bit roundings of the tensor-core path dominate the per-step eps error.

Runs the fp32 oracle with selected operands rounded to bf16 / fp16 (emulating operand storage of the
tcgen05 path: fp32 accumulation, fp32 residual stream) and reports eps rel-L2 vs the unrounded oracle.
Development aid only (imports oracle/); not part of the product or the tests.
"""
import os
import sys
import types

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ldm_oracle as O  # noqa: E402


def rnd(kind):
    if kind == "bf16":
        return lambda t: t.bfloat16().float()
    if kind == "fp16":
        return lambda t: t.half().float()
    return lambda t: t


class Hooks:
    def __init__(self, w="fp32", a_conv="fp32", a_lin="fp32", qkv="fp32", p="fp32", o="fp32"):
        self.w, self.a_conv, self.a_lin, self.qkv, self.p, self.o = map(rnd, (w, a_conv, a_lin, qkv, p, o))


def run(cfg, sd, x, t, ctx, hk: Hooks):
    Fp = types.SimpleNamespace(**{k: getattr(F, k) for k in dir(F) if not k.startswith("_")})

    def conv2d(inp, w, b=None, **kw):
        return F.conv2d(hk.a_conv(inp), hk.w(w), b, **kw)

    def linear(inp, w, b=None):
        if inp.dim() == 3 and inp.shape[1] > 1:
            return F.linear(hk.a_lin(inp), hk.w(w), b)
        return F.linear(inp, w, b)

    Fp.conv2d, Fp.linear = conv2d, linear

    def cross_attention(sd_, p, x_, context, heads, mask=None):
        if context is not None:
            return orig_ca(sd_, p, x_, context, heads, mask)
        b, n, c = x_.shape
        d = c // heads
        q = linear(x_, sd_[p + ".to_q.weight"])
        k = linear(x_, sd_[p + ".to_k.weight"])
        v = linear(x_, sd_[p + ".to_v.weight"])

        def split(t_):
            return t_.reshape(b, n, heads, d).permute(0, 2, 1, 3).reshape(b * heads, n, d)

        q, k, v = (hk.qkv(split(z)) for z in (q, k, v))
        sim = torch.einsum("bid,bjd->bij", q, k) * (d ** -0.5)
        m = sim.amax(-1, keepdim=True)
        pe = torch.exp(sim - m)
        l = pe.sum(-1, keepdim=True)   # fp32 row sum of unrounded p (as the kernel does)
        out = torch.einsum("bij,bjd->bid", hk.p(pe), v) / l
        out = hk.o(out)
        out = out.reshape(b, heads, n, d).permute(0, 2, 1, 3).reshape(b, n, c)
        return linear_noround_a(out, sd_[p + ".to_out.0.weight"], sd_[p + ".to_out.0.bias"])

    def linear_noround_a(inp, w, b):  # attention output is already rounded by hk.o
        return F.linear(inp, hk.w(w), b)

    orig_ca, orig_F = O.cross_attention, O.F
    O.cross_attention, O.F = cross_attention, Fp
    try:
        return O.unet_forward(sd, cfg, x, t, ctx)
    finally:
        O.cross_attention, O.F = orig_ca, orig_F


def main():
    hw = int(os.environ.get("HW", "64"))
    cfg = dict(O.FR_UNET)
    torch.manual_seed(0)
    sd = O.init_state_dict(cfg, 0)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(1, 3, hw, hw, generator=g)
    ctx = torch.randn(1, 1, 512, generator=g)
    t = torch.tensor([981])
    ref = run(cfg, sd, x, t, ctx, Hooks())
    cases = {
        "all bf16": dict(w="bf16", a_conv="bf16", a_lin="bf16", qkv="bf16", p="bf16", o="bf16"),
        "all fp16": dict(w="fp16", a_conv="fp16", a_lin="fp16", qkv="fp16", p="fp16", o="fp16"),
        "W bf16 only": dict(w="bf16"),
        "A(conv) bf16 only": dict(a_conv="bf16"),
        "A(lin) bf16 only": dict(a_lin="bf16"),
        "qkv bf16 only": dict(qkv="bf16"),
        "P bf16 only": dict(p="bf16"),
        "O bf16 only": dict(o="bf16"),
        "W bf16, rest fp16": dict(w="bf16", a_conv="fp16", a_lin="fp16", qkv="fp16", p="fp16", o="fp16"),
        "W fp16, A bf16": dict(w="fp16", a_conv="bf16", a_lin="bf16", qkv="bf16", p="bf16", o="bf16"),
    }
    for name, kw in cases.items():
        out = run(cfg, sd, x, t, ctx, Hooks(**kw))
        print(f"{name:24s} eps rel-L2 = {O.rel_l2(out, ref):.3e}", flush=True)


if __name__ == "__main__":
    main()
