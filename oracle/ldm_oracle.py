"""CPU oracle for the latent-diffusion denoising hot path (TEST INFRASTRUCTURE ONLY).

This file is a *restatement* of the reference algorithm — GiannisPikoulis/dsml-thesis,
``face_reenactment/ldm`` (byte-identical in ``talking_face/ldm`` for the UNet) —
as plain functional fp32 torch-CPU code driven by a flat ``state_dict``.  It is the checker
for the CUDA path, never the product: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  The package
``dsml_thesis_b200`` must never import anything under ``oracle/``.

Parity pinning: the reference ships no golden vectors or tests for this path, so
the oracle is pinned against the REAL reference modules imported from ``/root/reference`` in
the build container (``oracle/ref_shim.py`` + ``tests/golden/make_golden.py``); the resulting
vectors are committed under ``tests/golden/`` and re-checked on every CPU test run.

Every function cites the reference file:line it restates (paths relative to
``/root/reference/face_reenactment/``).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# ----------------------------------------------------------------------------------------
# configs of the two shipped UNets (configs/latent-diffusion/*.yaml)
# ----------------------------------------------------------------------------------------

#: face_reenactment/configs/latent-diffusion/affectnet-128-ldm-vq-f4.yaml:19-41
FR_UNET = dict(image_size=32, in_channels=3, out_channels=3, model_channels=160,
               attention_resolutions=[4, 2, 1], num_res_blocks=2, channel_mult=[1, 2, 4],
               num_head_channels=32, use_spatial_transformer=True, transformer_depth=1,
               context_dim=512)
#: talking_face/configs/latent-diffusion/mead-128-ldm-f4.yaml:20-42
TF_UNET = dict(FR_UNET, in_channels=9, context_dim=1024)
#: a small UNet with the same topology classes (channel counts stay multiples of 32, d_head 32)
TINY_UNET = dict(image_size=8, in_channels=3, out_channels=3, model_channels=32,
                 attention_resolutions=[4, 2, 1], num_res_blocks=1, channel_mult=[1, 2, 4],
                 num_head_channels=32, use_spatial_transformer=True, transformer_depth=1,
                 context_dim=64)
#: schedule constants, affectnet-128-ldm-vq-f4.yaml:5-9
LINEAR_START, LINEAR_END, NUM_TIMESTEPS = 0.0015, 0.0205, 1000


# ----------------------------------------------------------------------------------------
# UNet structure (openaimodel.py:516-686) -> list of layer records with state_dict prefixes
# ----------------------------------------------------------------------------------------

def unet_structure(cfg: dict) -> dict:
    """Mirror of ``UNetModel.__init__`` (openaimodel.py:506-686) for ``use_spatial_transformer=True``,
    ``resblock_updown=False``, ``conv_resample=True``, ``use_scale_shift_norm=False``.

    Returns ``{"input": [[layer,...],...], "middle": [...], "output": [[...],...]}`` where a
    layer is ``("conv", prefix, cin, cout)``, ``("res", prefix, cin, cout)``,
    ``("st", prefix, ch, heads, d_head)``, ``("down", prefix, ch)`` or ``("up", prefix, ch)``.
    """
    mc = cfg["model_channels"]
    mult = list(cfg["channel_mult"])
    nrb = cfg["num_res_blocks"]
    att = set(cfg["attention_resolutions"])
    nhc = cfg["num_head_channels"]
    inp: List[list] = [[("conv", "input_blocks.0.0", cfg["in_channels"], mc)]]
    chans = [mc]
    ch, ds = mc, 1
    for level, m in enumerate(mult):
        for _ in range(nrb):
            i = len(inp)
            layers = [("res", f"input_blocks.{i}.0", ch, m * mc)]
            ch = m * mc
            if ds in att:
                layers.append(("st", f"input_blocks.{i}.1", ch, ch // nhc, nhc))
            inp.append(layers)
            chans.append(ch)
        if level != len(mult) - 1:
            i = len(inp)
            inp.append([("down", f"input_blocks.{i}.0", ch)])
            chans.append(ch)
            ds *= 2
    mid = [("res", "middle_block.0", ch, ch), ("st", "middle_block.1", ch, ch // nhc, nhc),
           ("res", "middle_block.2", ch, ch)]
    out: List[list] = []
    for level, m in list(enumerate(mult))[::-1]:
        for i in range(nrb + 1):
            ich = chans.pop()
            o = len(out)
            layers = [("res", f"output_blocks.{o}.0", ch + ich, mc * m)]
            ch = mc * m
            if ds in att:
                layers.append(("st", f"output_blocks.{o}.{len(layers)}", ch, ch // nhc, nhc))
            if level and i == nrb:
                layers.append(("up", f"output_blocks.{o}.{len(layers)}", ch))
                ds //= 2
            out.append(layers)
    return {"input": inp, "middle": mid, "output": out, "out_ch": ch}


def param_shapes(cfg: dict) -> Dict[str, Tuple[int, ...]]:
    """Names and shapes of the UNet's parameters (626 tensors for FR/TF)."""
    mc, ctx = cfg["model_channels"], cfg["context_dim"]
    ted = 4 * mc
    s: Dict[str, Tuple[int, ...]] = {}

    def lin(p, i, o, bias=True):
        s[p + ".weight"] = (o, i)
        if bias:
            s[p + ".bias"] = (o,)

    def conv(p, i, o, k):
        s[p + ".weight"] = (o, i, k, k)
        s[p + ".bias"] = (o,)

    def norm(p, c):
        s[p + ".weight"] = (c,)
        s[p + ".bias"] = (c,)

    lin("time_embed.0", mc, ted)
    lin("time_embed.2", ted, ted)
    st = unet_structure(cfg)
    for layer in [l for blk in st["input"] for l in blk] + st["middle"] + [l for blk in st["output"] for l in blk]:
        kind, p = layer[0], layer[1]
        if kind == "conv":
            conv(p, layer[2], layer[3], 3)
        elif kind == "res":
            cin, cout = layer[2], layer[3]
            norm(p + ".in_layers.0", cin)
            conv(p + ".in_layers.2", cin, cout, 3)
            lin(p + ".emb_layers.1", ted, cout)
            norm(p + ".out_layers.0", cout)
            conv(p + ".out_layers.3", cout, cout, 3)
            if cin != cout:
                conv(p + ".skip_connection", cin, cout, 1)
        elif kind == "st":
            c = layer[2]
            norm(p + ".norm", c)
            conv(p + ".proj_in", c, c, 1)
            t = p + ".transformer_blocks.0"
            for a, cd in (("attn1", c), ("attn2", ctx)):
                lin(f"{t}.{a}.to_q", c, c, bias=False)
                lin(f"{t}.{a}.to_k", cd, c, bias=False)
                lin(f"{t}.{a}.to_v", cd, c, bias=False)
                lin(f"{t}.{a}.to_out.0", c, c)
            lin(f"{t}.ff.net.0.proj", c, 8 * c)
            lin(f"{t}.ff.net.2", 4 * c, c)
            for n in ("norm1", "norm2", "norm3"):
                norm(f"{t}.{n}", c)
            conv(p + ".proj_out", c, c, 1)
        elif kind == "down":
            conv(p + ".op", layer[2], layer[2], 3)
        elif kind == "up":
            conv(p + ".conv", layer[2], layer[2], 3)
    norm("out.0", st["out_ch"])
    conv("out.2", mc, cfg["out_channels"], 3)
    return s


def init_state_dict(cfg: dict, seed: int = 0) -> Dict[str, Tensor]:
    """Deterministic synthetic weights shared by reference, oracle and product in the tests.

    A freshly built reference UNet outputs exactly 0 (34 ``zero_module`` layers), so
    parity on it is vacuous; here EVERY tensor is non-zero: weights ~ U(-1,1)/sqrt(fan_in)
    (the ``kaiming_uniform(a=sqrt(5))`` bound torch uses), biases ~ 0.05·N(0,1), norm gains
    ~ 1+0.1·N(0,1).  Sorted-name order with one generator makes it independent of module
    construction order.
    """
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    for name, shape in sorted(param_shapes(cfg).items()):
        if len(shape) >= 2:
            fan_in = int(np.prod(shape[1:]))
            w = (torch.rand(shape, generator=g) * 2 - 1) / math.sqrt(fan_in)
            # sqrt(3) restores unit gain per layer so activations stay O(1) through the net
            sd[name] = (w * math.sqrt(3.0)).contiguous()
        elif "norm" in name or name.startswith("out.0") or ".in_layers.0." in name or ".out_layers.0." in name:
            if name.endswith("weight"):
                sd[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
            else:
                sd[name] = 0.1 * torch.randn(shape, generator=g)
        else:
            sd[name] = 0.05 * torch.randn(shape, generator=g)
    return sd


def state_dict_checksum(sd: Dict[str, Tensor]) -> Tuple[float, float]:
    """(Σx, Σ|x|) in float64 over all tensors in sorted-name order — detects RNG drift between boxes."""
    a = b = 0.0
    for k in sorted(sd):
        t = sd[k].double()
        a += float(t.sum())
        b += float(t.abs().sum())
    return a, b


# ----------------------------------------------------------------------------------------
# UNet forward (openaimodel.py:710-742)
# ----------------------------------------------------------------------------------------

def timestep_embedding(timesteps: Tensor, dim: int, max_period: int = 10000) -> Tensor:
    """util.py:151-171 — cos first, then sin; fp32 frequencies; ``timesteps.float()``."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32, device=timesteps.device) / half)
    args = timesteps[:, None].float() * freqs[None]
    emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        emb = torch.cat([emb, torch.zeros_like(emb[:, :1])], dim=-1)
    return emb


def group_norm32(x: Tensor, w: Tensor, b: Tensor, eps: float) -> Tensor:
    """util.py:214-216 (GroupNorm32, eps 1e-5) / attention.py:76-77 (Normalize, eps 1e-6); 32 groups."""
    return F.group_norm(x.float(), 32, w, b, eps).type(x.dtype)


def res_block(sd, p: str, x: Tensor, emb: Tensor) -> Tensor:
    """openaimodel.py:255-275 (``_forward``, no up/down, no scale-shift)."""
    h = F.conv2d(F.silu(group_norm32(x, sd[p + ".in_layers.0.weight"], sd[p + ".in_layers.0.bias"], 1e-5)),
                 sd[p + ".in_layers.2.weight"], sd[p + ".in_layers.2.bias"], padding=1)
    emb_out = F.linear(F.silu(emb), sd[p + ".emb_layers.1.weight"], sd[p + ".emb_layers.1.bias"])
    h = h + emb_out[..., None, None]
    h = F.conv2d(F.silu(group_norm32(h, sd[p + ".out_layers.0.weight"], sd[p + ".out_layers.0.bias"], 1e-5)),
                 sd[p + ".out_layers.3.weight"], sd[p + ".out_layers.3.bias"], padding=1)
    if p + ".skip_connection.weight" in sd:
        x = F.conv2d(x, sd[p + ".skip_connection.weight"], sd[p + ".skip_connection.bias"])
    return x + h


def cross_attention(sd, p: str, x: Tensor, context: Optional[Tensor], heads: int,
                    mask: Optional[Tensor] = None) -> Tensor:
    """attention.py:170-193 — bias-free q/k/v, scale on the logits, softmax over keys, to_out with bias."""
    b, n, c = x.shape
    d = c // heads
    ctx = x if context is None else context
    q = F.linear(x, sd[p + ".to_q.weight"])
    k = F.linear(ctx, sd[p + ".to_k.weight"])
    v = F.linear(ctx, sd[p + ".to_v.weight"])

    def split(t):  # 'b n (h d) -> (b h) n d' : head index is the OUTER factor (attention.py:178)
        return t.reshape(b, t.shape[1], heads, d).permute(0, 2, 1, 3).reshape(b * heads, t.shape[1], d)

    q, k, v = split(q), split(k), split(v)
    sim = torch.einsum("bid,bjd->bij", q, k) * (d ** -0.5)
    if mask is not None:  # attention.py:182-186
        m = mask.reshape(b, -1)[:, None, None, :].expand(b, heads, 1, -1).reshape(b * heads, 1, -1)
        sim = sim.masked_fill(~m, -torch.finfo(sim.dtype).max)
    attn = sim.softmax(dim=-1)
    out = torch.einsum("bij,bjd->bid", attn, v)
    out = out.reshape(b, heads, n, d).permute(0, 2, 1, 3).reshape(b, n, c)
    return F.linear(out, sd[p + ".to_out.0.weight"], sd[p + ".to_out.0.bias"])


def basic_transformer_block(sd, p: str, x: Tensor, context: Optional[Tensor], heads: int) -> Tensor:
    """attention.py:211-215 (+ GEGLU 43-44: first half value, second half gate, exact-erf gelu)."""
    c = x.shape[-1]

    def ln(n, t):
        return F.layer_norm(t, (c,), sd[f"{p}.{n}.weight"], sd[f"{p}.{n}.bias"], 1e-5)

    x = cross_attention(sd, p + ".attn1", ln("norm1", x), None, heads) + x
    x = cross_attention(sd, p + ".attn2", ln("norm2", x), context, heads) + x
    hcat = F.linear(ln("norm3", x), sd[p + ".ff.net.0.proj.weight"], sd[p + ".ff.net.0.proj.bias"])
    a, gate = hcat.chunk(2, dim=-1)
    x = F.linear(a * F.gelu(gate), sd[p + ".ff.net.2.weight"], sd[p + ".ff.net.2.bias"]) + x
    return x


def spatial_transformer(sd, p: str, x: Tensor, context: Optional[Tensor], heads: int) -> Tensor:
    """attention.py:250-261."""
    b, c, h, w = x.shape
    x_in = x
    x = F.group_norm(x, 32, sd[p + ".norm.weight"], sd[p + ".norm.bias"], 1e-6)
    x = F.conv2d(x, sd[p + ".proj_in.weight"], sd[p + ".proj_in.bias"])
    x = x.permute(0, 2, 3, 1).reshape(b, h * w, c)
    x = basic_transformer_block(sd, p + ".transformer_blocks.0", x, context, heads)
    x = x.reshape(b, h, w, c).permute(0, 3, 1, 2)
    x = F.conv2d(x, sd[p + ".proj_out.weight"], sd[p + ".proj_out.bias"])
    return x + x_in


def run_layers(sd, layers, h: Tensor, emb: Tensor, context) -> Tensor:
    """TimestepEmbedSequential.forward, openaimodel.py:80-88."""
    for layer in layers:
        kind, p = layer[0], layer[1]
        if kind == "conv":
            h = F.conv2d(h, sd[p + ".weight"], sd[p + ".bias"], padding=1)
        elif kind == "res":
            h = res_block(sd, p, h, emb)
        elif kind == "st":
            h = spatial_transformer(sd, p, h, context, layer[3])
        elif kind == "down":  # openaimodel.py:151-160
            h = F.conv2d(h, sd[p + ".op.weight"], sd[p + ".op.bias"], stride=2, padding=1)
        elif kind == "up":  # openaimodel.py:109-119
            h = F.interpolate(h, scale_factor=2, mode="nearest")
            h = F.conv2d(h, sd[p + ".conv.weight"], sd[p + ".conv.bias"], padding=1)
    return h


def unet_forward_autograd(sd: Dict[str, Tensor], cfg: dict, x: Tensor, timesteps: Tensor,
                          context: Optional[Tensor] = None, taps: Optional[dict] = None) -> Tensor:
    """openaimodel.py:710-742 with autograd left on (the training oracle differentiates through it).
    ``taps`` (optional dict) receives each block's output for block-level parity."""
    st = unet_structure(cfg)
    t_emb = timestep_embedding(timesteps, cfg["model_channels"])
    emb = F.linear(F.silu(F.linear(t_emb, sd["time_embed.0.weight"], sd["time_embed.0.bias"])),
                   sd["time_embed.2.weight"], sd["time_embed.2.bias"])
    hs = []
    h = x.float()
    for i, blk in enumerate(st["input"]):
        h = run_layers(sd, blk, h, emb, context)
        hs.append(h)
        if taps is not None:
            taps[f"input_blocks.{i}"] = h
    h = run_layers(sd, st["middle"], h, emb, context)
    if taps is not None:
        taps["middle_block"] = h
    for i, blk in enumerate(st["output"]):
        h = torch.cat([h, hs.pop()], dim=1)
        h = run_layers(sd, blk, h, emb, context)
        if taps is not None:
            taps[f"output_blocks.{i}"] = h
    h = h.type(x.dtype)
    h = F.silu(group_norm32(h, sd["out.0.weight"], sd["out.0.bias"], 1e-5))
    return F.conv2d(h, sd["out.2.weight"], sd["out.2.bias"], padding=1)


@torch.no_grad()
def unet_forward(sd: Dict[str, Tensor], cfg: dict, x: Tensor, timesteps: Tensor,
                 context: Optional[Tensor] = None, taps: Optional[dict] = None) -> Tensor:
    """openaimodel.py:710-742 (inference: no autograd graph)."""
    return unet_forward_autograd(sd, cfg, x, timesteps, context, taps)


# ----------------------------------------------------------------------------------------
# training step: LatentDiffusion.p_losses (ddpm.py:1014-1047; talking_face ddpm2cond.py:965-998)
# ----------------------------------------------------------------------------------------

def q_sample(x_start: Tensor, t: Tensor, noise: Tensor, bufs: Optional[dict] = None) -> Tensor:
    """ddpm.py:273-276: extract_into_tensor(sqrt_alphas_cumprod, t) * x_start + extract(sqrt_one_minus, t) * noise."""
    bufs = ddpm_buffers() if bufs is None else bufs
    e = lambda a: a.gather(-1, t).reshape(-1, 1, 1, 1)
    return e(bufs["sqrt_alphas_cumprod"]) * x_start + e(bufs["sqrt_one_minus_alphas_cumprod"]) * noise


def p_losses(sd: Dict[str, Tensor], cfg: dict, x_start: Tensor, t: Tensor, noise: Tensor, context: Tensor,
             c_concat: Optional[Tensor] = None):
    """ddpm.py:1014-1047 for parameterization "eps", loss_type "l2", logvar == 0 (not learned), l_simple_weight 1,
    original_elbo_weight 0:  x_noisy = q_sample; eps = apply_model (TF: cat([x_noisy, c34], 1), ddpm2cond.py:1307-1311);
    loss_simple = mse(eps, noise).mean([1,2,3]); loss = loss_simple.mean().
    Returns ``(loss, grads)`` with ``grads`` = d loss / d sd[name] for every parameter, plus ``"context"``."""
    leaves = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    ctx = context.detach().clone().requires_grad_(True)
    x_noisy = q_sample(x_start, t, noise)
    x_in = torch.cat([x_noisy, c_concat], dim=1) if c_concat is not None else x_noisy
    eps = unet_forward_autograd(leaves, cfg, x_in, t, ctx)
    loss_simple = ((eps - noise) ** 2).mean([1, 2, 3])
    loss = loss_simple.mean()
    names = list(leaves)
    gs = torch.autograd.grad(loss, [leaves[k] for k in names] + [ctx], allow_unused=True)
    grads = {k: (g if g is not None else torch.zeros_like(leaves[k])) for k, g in zip(names, gs[:-1])}
    grads["context"] = gs[-1] if gs[-1] is not None else torch.zeros_like(ctx)
    return loss.detach(), grads


# ----------------------------------------------------------------------------------------
# schedules (util.py:21-74, ddpm.py:117-133, ddim.py:24-53)
# ----------------------------------------------------------------------------------------

def make_beta_schedule_linear(n_timestep=NUM_TIMESTEPS, linear_start=LINEAR_START, linear_end=LINEAR_END):
    """util.py:22-25 — ``linspace(sqrt(s), sqrt(e), T)**2`` in float64."""
    return (torch.linspace(linear_start ** 0.5, linear_end ** 0.5, n_timestep, dtype=torch.float64) ** 2).numpy()


def ddpm_buffers(n_timestep=NUM_TIMESTEPS, linear_start=LINEAR_START, linear_end=LINEAR_END):
    """ddpm.py:117-142 — fp32 buffers built from float64 numpy."""
    betas = make_beta_schedule_linear(n_timestep, linear_start, linear_end)
    ac = np.cumprod(1.0 - betas, axis=0)
    ac_prev = np.append(1.0, ac[:-1])
    f32 = lambda a: torch.tensor(a, dtype=torch.float32)
    return dict(betas=f32(betas), alphas_cumprod=f32(ac), alphas_cumprod_prev=f32(ac_prev),
                sqrt_alphas_cumprod=f32(np.sqrt(ac)), sqrt_one_minus_alphas_cumprod=f32(np.sqrt(1.0 - ac)))


def make_ddim_timesteps(num_ddim: int, num_ddpm: int = NUM_TIMESTEPS, method: str = "uniform") -> np.ndarray:
    """util.py:46-60."""
    if method == "uniform":
        c = num_ddpm // num_ddim
        ts = np.asarray(list(range(0, num_ddpm, c)))
    elif method == "quad":
        ts = ((np.linspace(0, np.sqrt(num_ddpm * .8), num_ddim)) ** 2).astype(int)
    else:
        raise NotImplementedError(method)
    return ts + 1


def ddim_schedule(alphas_cumprod_f32: Tensor, num_ddim: int, eta: float, method: str = "uniform") -> dict:
    """ddim.py:24-53 + util.py:63-74.  Returns the four per-step coefficient arrays exactly as
    ``p_sample_ddim`` consumes them: each value rounded to fp32 the way ``torch.full`` rounds it
    (SURVEY Appendix B), plus the timesteps."""
    ts = make_ddim_timesteps(num_ddim, alphas_cumprod_f32.shape[0], method)
    ac = alphas_cumprod_f32.cpu()
    alphas = ac[ts]                                                    # fp32 tensor
    alphas_prev = np.asarray([ac[0]] + ac[ts[:-1]].tolist())           # float64 numpy of fp32 values
    # sigmas = eta * np.sqrt((1 - alphas_prev) / (1 - alphas) * (1 - alphas / alphas_prev))   (util.py:69)
    # dtype path of that literal expression [verified against the live reference]: ``alphas`` is a torch
    # fp32 tensor and ``alphas_prev`` float64 numpy, so ``ndarray / Tensor`` dispatches to
    # ``Tensor.__rtruediv__`` = ``(1 - alphas).reciprocal()`` IN FP32, then times the float64 numerator;
    # ``alphas / alphas_prev`` is a true float64 division; sqrt in float64.
    ap64 = torch.from_numpy(alphas_prev)
    sig = eta * torch.sqrt((1 - alphas).reciprocal() * (1 - ap64) * (1 - alphas / ap64))
    sqrt_1ma = np.sqrt((1.0 - alphas).numpy())                         # fp32 subtract in torch, numpy sqrt (ddim.py:49)
    return dict(timesteps=ts,
                a_t=alphas.numpy().astype(np.float32),
                a_prev=alphas_prev.astype(np.float32),
                sigma_t=sig.numpy().astype(np.float32),
                sqrt_one_minus_at=np.asarray(sqrt_1ma, dtype=np.float32))


def ddim_update(x: Tensor, e_t: Tensor, a_t, a_prev, sigma_t, sqrt_one_minus_at,
                noise: Optional[Tensor] = None, temperature: float = 1.0) -> Tuple[Tensor, Tensor]:
    """ddim.py:188-203 — all coefficient arithmetic in fp32 on ``(b,1,1,1)`` tensors."""
    b = x.shape[0]
    full = lambda v: torch.full((b, 1, 1, 1), float(v), dtype=torch.float32)
    a_t, a_prev, sigma_t, s1 = full(a_t), full(a_prev), full(sigma_t), full(sqrt_one_minus_at)
    pred_x0 = (x - s1 * e_t) / a_t.sqrt()
    dir_xt = (1. - a_prev - sigma_t ** 2).sqrt() * e_t
    nz = sigma_t * (noise if noise is not None else torch.zeros_like(x)) * temperature
    x_prev = a_prev.sqrt() * pred_x0 + dir_xt + nz
    return x_prev, pred_x0


@torch.no_grad()
def ddim_sample(apply_model, alphas_cumprod_f32: Tensor, S: int, x_T: Tensor, cond, eta: float = 0.0,
                unconditional_guidance_scale: float = 1.0, unconditional_conditioning=None,
                noises: Optional[Sequence[Tensor]] = None, temperature: float = 1.0,
                log_every_t: int = 100):
    """ddim.py:113-203 (``ddim_sampling`` + ``p_sample_ddim``), FR flavour: CFG by batch doubling in
    ``[uncond, cond]`` order (173-177).  ``noises[i]`` is the N(0,1) draw for loop iteration ``i``
    (the reference calls ``torch.randn`` per step, util.py:264-267)."""
    sch = ddim_schedule(alphas_cumprod_f32, S, eta)
    ts = sch["timesteps"]
    total = ts.shape[0]
    img = x_T
    b = x_T.shape[0]
    inter = {"x_inter": [img], "pred_x0": [img]}
    eps_trace = []
    for i, step in enumerate(np.flip(ts)):
        index = total - i - 1
        t = torch.full((b,), int(step), dtype=torch.long)
        if unconditional_conditioning is None or unconditional_guidance_scale == 1.0:
            e_t = apply_model(img, t, cond)
        else:
            e_u, e_c = apply_model(torch.cat([img] * 2), torch.cat([t] * 2),
                                   torch.cat([unconditional_conditioning, cond])).chunk(2)
            e_t = e_u + unconditional_guidance_scale * (e_c - e_u)
        eps_trace.append(e_t)
        nz = noises[i] if (noises is not None and eta > 0) else None
        img, pred_x0 = ddim_update(img, e_t, sch["a_t"][index], sch["a_prev"][index], sch["sigma_t"][index],
                                   sch["sqrt_one_minus_at"][index], nz, temperature)
        if index % log_every_t == 0 or index == total - 1:
            inter["x_inter"].append(img)
            inter["pred_x0"].append(pred_x0)
    return img, inter, eps_trace


def make_ddim_timesteps_strength(num_ddim: int, num_ddpm: int = NUM_TIMESTEPS, strength: float = 1.0) -> np.ndarray:
    """ddim2.py:9-29 / latent_manipulation.py:52-72 ('uniform'): linspace over [0, int(T*strength)], truncated
    to int, first entry forced to 1, no +1 shift."""
    ts = np.linspace(0, 1, num_ddim) * int(num_ddpm * strength)
    ts = [int(v) for v in list(ts)]
    return np.asarray([1] + ts[1:])


def ddim_schedule_strength(alphas_cumprod_f32: Tensor, num_ddim: int, strength: float, eta: float = 0.0) -> dict:
    """ddim2.py:66-97 (DDIMSampler2.make_schedule with ``strength``): the coefficient arrays of ddim_schedule on the
    strength-scaled grid plus ``sqrt_one_minus_a_prev`` = np.sqrt(1. - ddim_alphas_prev) (float64 numpy, ddim2.py:96),
    every value rounded to fp32 the way ``torch.full`` rounds it."""
    ts = make_ddim_timesteps_strength(num_ddim, alphas_cumprod_f32.shape[0], strength)
    ac = alphas_cumprod_f32.cpu()
    alphas = ac[ts]
    alphas_prev = np.asarray([ac[0]] + ac[ts[:-1]].tolist())
    ap64 = torch.from_numpy(alphas_prev)
    sig = eta * torch.sqrt((1 - alphas).reciprocal() * (1 - ap64) * (1 - alphas / ap64))
    sqrt_1ma = np.sqrt((1.0 - alphas).numpy())
    return dict(timesteps=ts,
                a_t=alphas.numpy().astype(np.float32),
                a_prev=alphas_prev.astype(np.float32),
                sigma_t=sig.numpy().astype(np.float32),
                sqrt_one_minus_at=np.asarray(sqrt_1ma, dtype=np.float32),
                sqrt_one_minus_a_prev=np.sqrt(1. - alphas_prev).astype(np.float32))


def ddim_invert_update(x: Tensor, e_t: Tensor, a_t, a_next, sqrt_one_minus_at) -> Tuple[Tensor, Tensor]:
    """latent_manipulation.py:404-419 (``q_sample_ddim``): x_t -> x_{t+1}.  ``a_t`` = ddim_alphas_prev[index],
    ``a_next`` = ddim_alphas[index], ``sqrt_one_minus_at`` = ddim_sqrt_one_minus_alphas_prev[index]."""
    b = x.shape[0]
    full = lambda v: torch.full((b, 1, 1, 1), float(v), dtype=torch.float32)
    at, at_next, s1 = full(a_t), full(a_next), full(sqrt_one_minus_at)
    pred_x0 = (x - s1 * e_t) / at.sqrt()
    dir_xt = (1. - at_next).sqrt() * e_t
    return at_next.sqrt() * pred_x0 + dir_xt, pred_x0


@torch.no_grad()
def ddim_latent_manipulation(apply_model, alphas_cumprod_f32: Tensor, S: int, x0: Tensor, c_src, c_trg,
                             strength: float = 1.0, unconditional_guidance_scale: float = 1.0,
                             unconditional_conditioning=None):
    """latent_manipulation.py:421-489: DDIM inversion of ``x0`` under ``c_src`` over the strength-scaled grid, then
    deterministic (eta = 0) re-generation under ``c_trg``.  Returns ``(img, x_inverted, x0)``."""
    sch = ddim_schedule_strength(alphas_cumprod_f32, S, strength, 0.0)
    ts = sch["timesteps"]
    total = ts.shape[0]
    b = x0.shape[0]

    def eps(img, t, cond):
        if unconditional_conditioning is None or unconditional_guidance_scale == 1.0:
            return apply_model(img, t, cond)
        e_u, e_c = apply_model(torch.cat([img] * 2), torch.cat([t] * 2),
                               torch.cat([unconditional_conditioning, cond])).chunk(2)
        return e_u + unconditional_guidance_scale * (e_c - e_u)

    x = x0.clone()
    for i, step in enumerate(ts):
        t = torch.full((b,), int(step), dtype=torch.long)
        x, _ = ddim_invert_update(x, eps(x, t, c_src), sch["a_prev"][i], sch["a_t"][i], sch["sqrt_one_minus_a_prev"][i])
    img = x.clone()
    for i, step in enumerate(np.flip(ts)):
        index = total - i - 1
        t = torch.full((b,), int(step), dtype=torch.long)
        img, _ = ddim_update(img, eps(img, t, c_trg), sch["a_t"][index], sch["a_prev"][index], sch["sigma_t"][index],
                             sch["sqrt_one_minus_at"][index])
    return img, x, x0


def rel_l2(a: Tensor, b: Tensor) -> float:
    """‖a−b‖₂ / ‖b‖₂ in float64 — the parity metric of BASELINE.json's north_star."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))
