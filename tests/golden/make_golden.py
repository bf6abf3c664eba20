"""Generate the committed golden vectors from the REAL reference (run in the build container only).

    python tests/golden/make_golden.py        # writes tests/golden/*.npz

The reference ships no tests or golden vectors for this path, so these files are what
pins the oracle: every array here is an output of the unmodified reference modules
(``/root/reference/{face_reenactment,talking_face}/ldm``: ``UNetModel``, ``DDIMSampler``,
``timestep_embedding``, ``make_beta_schedule``) run on CPU fp32 with the synthetic weights of
``oracle.ldm_oracle.init_state_dict`` (loaded with ``strict=True`` — which also proves the 626
parameter names/shapes).  ``/root/reference`` does not exist on the GPU box; the tests there read
only the ``.npz`` files.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))

from oracle import ldm_oracle as O  # noqa: E402
from oracle import ref_shim as R  # noqa: E402


def synth_inputs(cfg, B, hw, seed, ctx_len=1):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, cfg["in_channels"], hw, hw, generator=g)
    ctx = torch.randn(B, ctx_len, cfg["context_dim"], generator=g)
    return x, ctx


def unet_goldens():
    out = {}
    cases = [  # name, cfg, flavour, B, hw, ctx_len, timesteps
        ("tiny_c1", O.TINY_UNET, "face_reenactment", 2, 8, 1, [981, 501, 21]),
        ("tiny_c3", O.TINY_UNET, "face_reenactment", 2, 8, 3, [981, 21]),
        ("tiny_hw16", O.TINY_UNET, "face_reenactment", 3, 16, 1, [501]),
        ("fr_32", O.FR_UNET, "face_reenactment", 1, 32, 1, [981]),
        ("tf_32", O.TF_UNET, "talking_face", 1, 32, 1, [501]),
    ]
    for name, cfg, flavour, B, hw, cl, tsteps in cases:
        sd = O.init_state_dict(cfg, seed=0)
        ref = R.reference_unet(cfg, flavour).eval()
        ref.load_state_dict(sd, strict=True)
        assert len(list(ref.buffers())) == 0
        x, ctx = synth_inputs(cfg, B, hw, seed=1, ctx_len=cl)
        out[f"{name}/checksum"] = np.asarray(O.state_dict_checksum(sd), dtype=np.float64)
        out[f"{name}/n_params"] = np.asarray([len(sd), sum(v.numel() for v in sd.values())], dtype=np.int64)
        for t in tsteps:
            tt = torch.tensor([t] * B, dtype=torch.long)
            with torch.no_grad():
                eps = ref(x, tt, context=ctx)
            out[f"{name}/eps_t{t}"] = eps.numpy()
        # per-sample distinct timesteps (training-style call, ddpm.py:1015)
        tt = torch.tensor([(37 * (i + 1)) % 1000 for i in range(B)], dtype=torch.long)
        with torch.no_grad():
            out[f"{name}/eps_tmixed"] = ref(x, tt, context=ctx).numpy()
        out[f"{name}/tmixed"] = tt.numpy()
        if name == "tiny_c1":
            # block-level taps through forward hooks on the reference module
            taps = {}
            hooks = []
            for bn, mod in [(f"input_blocks.{i}", m) for i, m in enumerate(ref.input_blocks)] + \
                           [("middle_block", ref.middle_block)] + \
                           [(f"output_blocks.{i}", m) for i, m in enumerate(ref.output_blocks)]:
                hooks.append(mod.register_forward_hook(lambda m, a, o, bn=bn: taps.__setitem__(bn, o.detach())))
            with torch.no_grad():
                ref(x, torch.tensor([981] * B), context=ctx)
            for h in hooks:
                h.remove()
            for bn, v in taps.items():
                out[f"{name}/tap/{bn}"] = v.numpy()
    return out


def unet64_goldens():
    """BASELINE configs 1-3 run 64x64 latents: one eval of the real FR / TF UNets at that size (B=1)."""
    out = {}
    for name, cfg, flavour, B, hw, cl, tsteps in [("fr_64", O.FR_UNET, "face_reenactment", 1, 64, 1, [981, 21]),
                                                  ("tf_64", O.TF_UNET, "talking_face", 1, 64, 1, [501])]:
        sd = O.init_state_dict(cfg, seed=0)
        ref = R.reference_unet(cfg, flavour).eval()
        ref.load_state_dict(sd, strict=True)
        x, ctx = synth_inputs(cfg, B, hw, seed=1, ctx_len=cl)
        for t in tsteps:
            with torch.no_grad():
                out[f"{name}/eps_t{t}"] = ref(x, torch.tensor([t] * B, dtype=torch.long), context=ctx).numpy()
    return out


def unet128_goldens():
    """BASELINE config 5 (roofline stress): the shipped FR architecture at 128x128 latents — attention over
    16384 / 4096 / 1024 tokens (SURVEY 8d reading C5a) — one eval of the real reference UNet, B=1."""
    out = {}
    cfg = O.FR_UNET
    sd = O.init_state_dict(cfg, seed=0)
    ref = R.reference_unet(cfg, "face_reenactment").eval()
    ref.load_state_dict(sd, strict=True)
    x, ctx = synth_inputs(cfg, 1, 128, seed=1, ctx_len=1)
    with torch.no_grad():
        out["fr_128/eps_t501"] = ref(x, torch.tensor([501], dtype=torch.long), context=ctx).numpy()
    return out


def schedule_goldens():
    out = {}
    R.import_reference("face_reenactment")
    from ldm.modules.diffusionmodules.util import timestep_embedding
    unet = torch.nn.Identity()
    model = R.LatentDiffusionStandIn(unet)
    out["ddpm/betas"] = model.betas.numpy()
    out["ddpm/alphas_cumprod"] = model.alphas_cumprod.numpy()
    out["ddpm/alphas_cumprod_prev"] = model.alphas_cumprod_prev.numpy()
    for S in (50, 200, 7):
        for eta in (0.0, 1.0):
            smp = R.cpu_ddim_sampler(model)
            smp.make_schedule(S, ddim_eta=eta, verbose=False)
            n = smp.ddim_timesteps.shape[0]
            # values exactly as p_sample_ddim materialises them: torch.full(..., v) in fp32 (ddim.py:188-191)
            f = lambda arr: np.asarray([torch.full((1,), arr[i]).item() for i in range(n)], dtype=np.float32)
            k = f"ddim/S{S}_eta{int(eta)}"
            out[k + "/timesteps"] = np.asarray(smp.ddim_timesteps, dtype=np.int64)
            out[k + "/a_t"] = f(smp.ddim_alphas)
            out[k + "/a_prev"] = f(smp.ddim_alphas_prev)
            out[k + "/sigma_t"] = f(smp.ddim_sigmas)
            out[k + "/sqrt_one_minus_at"] = f(smp.ddim_sqrt_one_minus_alphas)
    t = torch.tensor([0, 1, 21, 501, 981, 999])
    out["temb/t"] = t.numpy()
    out["temb/dim160"] = timestep_embedding(t, 160).numpy()
    out["temb/dim32"] = timestep_embedding(t, 32).numpy()
    return out


def trajectory_goldens():
    """Full DDIM trajectories of the reference sampler on the tiny UNet (seconds on CPU)."""
    out = {}
    cfg = O.TINY_UNET
    sd = O.init_state_dict(cfg, seed=0)
    B, hw = 2, 8
    # ---- FR flavour: ddim.py, eta 0 / eta 1 / CFG -------------------------------------------
    ref = R.reference_unet(cfg, "face_reenactment").eval()
    ref.load_state_dict(sd, strict=True)
    model = R.LatentDiffusionStandIn(ref)
    x_T, ctx = synth_inputs(cfg, B, hw, seed=2)
    g = torch.Generator().manual_seed(3)
    uc = torch.randn(B, 1, cfg["context_dim"], generator=g)
    for tag, S, eta, scale in [("fr_S10_eta0", 10, 0.0, 1.0), ("fr_S50_eta0", 50, 0.0, 1.0),
                               ("fr_S10_eta1", 10, 1.0, 1.0), ("fr_S10_cfg3", 10, 0.0, 3.0)]:
        smp = R.cpu_ddim_sampler(model)
        torch.manual_seed(1234)  # the per-step torch.randn draws of noise_like (util.py:264-267)
        samples, inter = smp.sample(S, batch_size=B, shape=[3, hw, hw], conditioning=ctx, eta=eta, x_T=x_T,
                                    verbose=False, log_every_t=3, unconditional_guidance_scale=scale,
                                    unconditional_conditioning=uc if scale != 1.0 else None)
        out[f"traj/{tag}/samples"] = samples.numpy()
        out[f"traj/{tag}/x_inter"] = torch.stack(inter["x_inter"]).numpy()
        out[f"traj/{tag}/pred_x0"] = torch.stack(inter["pred_x0"]).numpy()
    out["traj/x_T"] = x_T.numpy()
    out["traj/ctx"] = ctx.numpy()
    out["traj/uc"] = uc.numpy()
    # ---- TF flavour: ddim2cond.py with the 2-key dict conditioning --------------------------
    cfg9 = dict(cfg, in_channels=9)
    sd9 = O.init_state_dict(cfg9, seed=0)
    ref9 = R.reference_unet(cfg9, "talking_face").eval()
    ref9.load_state_dict(sd9, strict=True)
    model9 = R.LatentDiffusionStandIn(ref9, two_cond=True)
    g = torch.Generator().manual_seed(4)
    c34 = torch.randn(B, 6, hw, hw, generator=g)
    smp = R.cpu_ddim_sampler(model9, "talking_face")
    torch.manual_seed(1234)
    samples, x_inter = smp.sample(10, batch_size=B, shape=[3, hw, hw],
                                  conditioning={"class_label_&_audio": ctx, "motion_&_id": c34},
                                  eta=1.0, x_T=x_T, verbose=False, log_every_t=3)
    out["traj/tf_S10_eta1/samples"] = samples.numpy()
    out["traj/tf_S10_eta1/x_inter"] = torch.stack(x_inter).numpy()
    out["traj/c34"] = c34.numpy()
    return out


def manipulation_goldens():
    """DDIM inversion + re-generation from the reference's own ``DDIMSampler2`` schedule and the
    ``q_sample_ddim`` / ``latent_manipulation`` source of FR/latent_manipulation.py, tiny UNet on CPU."""
    out = {}
    cfg = O.TINY_UNET
    sd = O.init_state_dict(cfg, seed=0)
    B, hw = 2, 8
    ref = R.reference_unet(cfg, "face_reenactment").eval()
    ref.load_state_dict(sd, strict=True)
    model = R.LatentDiffusionStandIn(ref)
    x0, c_src = synth_inputs(cfg, B, hw, seed=5)
    g = torch.Generator().manual_seed(6)
    c_trg = torch.randn(B, 1, cfg["context_dim"], generator=g)
    uc = torch.randn(B, 1, cfg["context_dim"], generator=g)
    out["manip/x0"], out["manip/c_src"], out["manip/c_trg"], out["manip/uc"] = (t.numpy() for t in (x0, c_src, c_trg, uc))
    for S, strength in ((10, 0.8), (12, 0.6), (50, 0.35)):
        smp = R.cpu_manipulation_sampler(model)
        smp.make_schedule(S, ddim_eta=0.0, strength=strength, verbose=False)
        n = smp.ddim_timesteps.shape[0]
        f = lambda arr: np.asarray([torch.full((1,), arr[i]).item() for i in range(n)], dtype=np.float32)
        k = f"manip/S{S}_s{int(strength * 100)}"
        out[k + "/timesteps"] = np.asarray(smp.ddim_timesteps, dtype=np.int64)
        out[k + "/a_t"] = f(smp.ddim_alphas)
        out[k + "/a_prev"] = f(smp.ddim_alphas_prev)
        out[k + "/sqrt_one_minus_at"] = f(smp.ddim_sqrt_one_minus_alphas)
        out[k + "/sqrt_one_minus_a_prev"] = f(smp.ddim_sqrt_one_minus_alphas_prev)
    for tag, S, strength, scale in (("S10_s80", 10, 0.8, 1.0), ("S12_s60", 12, 0.6, 1.0), ("S8_s50_cfg2", 8, 0.5, 2.0)):
        smp = R.cpu_manipulation_sampler(model)
        import contextlib, io
        with contextlib.redirect_stdout(io.StringIO()):
            img, x_inv, _ = smp.latent_manipulation(c_src, c_trg, S, B, [3, hw, hw], x0=x0, eta=0., verbose=False,
                                                    strength=strength, unconditional_guidance_scale=scale,
                                                    unconditional_conditioning=uc if scale != 1.0 else None)
        out[f"manip/{tag}/img"] = img.numpy()
        out[f"manip/{tag}/x_inv"] = x_inv.numpy()
    return out


def train_goldens():
    """p_losses through the REAL reference UNetModel + torch autograd (ddpm.py:1014-1047 arithmetic spelled out with the
    reference's own q_sample / schedule; LatentDiffusion itself needs pytorch_lightning): loss and the gradient of
    a fingerprint (L2 norm + 64 / 8 evenly spaced entries) of the gradient of every parameter of the tiny / FR / TF UNets
    at 16x16 latents."""
    out = {}
    for tag, cfg, flavour, B, hw in (("tiny", O.TINY_UNET, "face_reenactment", 2, 16),
                                     ("fr16", O.FR_UNET, "face_reenactment", 2, 16),
                                     ("tf16", O.TF_UNET, "talking_face", 2, 16)):
        unet = R.reference_unet(cfg, flavour)
        sd = O.init_state_dict(cfg, 0)
        unet.load_state_dict(sd, strict=True)
        unet.train()
        model = R.LatentDiffusionStandIn(unet, two_cond=(flavour == "talking_face"))
        g = torch.Generator().manual_seed(77)
        x0 = torch.randn(B, 3, hw, hw, generator=g)
        noise = torch.randn(B, 3, hw, hw, generator=g)
        ctx = torch.randn(B, 1, cfg["context_dim"], generator=g).requires_grad_(True)
        c34 = torch.randn(B, 6, hw, hw, generator=g) if flavour == "talking_face" else None
        t = torch.tensor([731, 12][:B])
        x_noisy = model.q_sample(x0, t, noise)
        eps = model.apply_model(x_noisy, t, ctx, c34) if c34 is not None else model.apply_model(x_noisy, t, ctx)
        loss = torch.nn.functional.mse_loss(eps, noise, reduction="none").mean([1, 2, 3]).mean()
        loss.backward()
        out[f"train/{tag}/loss"] = loss.detach().numpy()
        out[f"train/{tag}/x_noisy"] = x_noisy.detach().numpy()
        out[f"train/{tag}/dctx"] = ctx.grad.numpy()
        for name, p in unet.named_parameters():
            gr = p.grad if p.grad is not None else torch.zeros_like(p)
            flat = gr.reshape(-1)
            idx = torch.linspace(0, flat.numel() - 1, 64 if tag == "tiny" else 8).long()
            out[f"train/{tag}/gradfp/{name}"] = torch.cat([gr.double().norm().reshape(1), flat[idx].double()]).numpy()
    return out


def vq_goldens():
    """First-stage decode through the REAL reference modules: ``VectorQuantizer2`` (taming, vendored in the
    reference tree), ``post_quant_conv`` and ``Decoder`` assembled exactly as ``VQModel.__init__`` does
    (autoencoder.py:36-43; ``VQModelInterface`` itself needs pytorch_lightning), synthetic weights of
    ``oracle.vq_oracle.init_vq_state_dict`` loaded with strict=True (which proves the parameter names / shapes)."""
    import contextlib
    import io
    from oracle import vq_oracle as V
    R.import_reference("face_reenactment")
    from ldm.modules.diffusionmodules.model import Decoder
    from taming.modules.vqvae.quantize import VectorQuantizer2

    class FirstStage(torch.nn.Module):
        def __init__(self, cfg):
            super().__init__()
            dd = cfg["ddconfig"]
            with contextlib.redirect_stdout(io.StringIO()):
                self.decoder = Decoder(**dd)
            self.quantize = VectorQuantizer2(cfg["n_embed"], cfg["embed_dim"], beta=0.25)
            self.post_quant_conv = torch.nn.Conv2d(cfg["embed_dim"], dd["z_channels"], 1)

        def decode(self, h, force_not_quantize=False):      # autoencoder.py:274-283 verbatim semantics
            info = None
            if not force_not_quantize:
                quant, _, info = self.quantize(h)
            else:
                quant = h
            return self.decoder(self.post_quant_conv(quant)), info

    out = {}
    for name, cfg, B, hw, seed in [("tiny_8", V.VQ_TINY, 2, 8, 5), ("tiny_16", V.VQ_TINY, 1, 16, 6),
                                   ("f4_16", V.VQ_F4, 1, 16, 7)]:
        sd = V.init_vq_state_dict(cfg, 0)
        m = FirstStage(cfg)
        m.load_state_dict(sd, strict=True)
        m.eval()
        g = torch.Generator().manual_seed(seed)
        z = torch.randn(B, cfg["embed_dim"], hw, hw, generator=g)
        with torch.no_grad():
            img_q, info = m.decode(z)
            img_n, _ = m.decode(z, force_not_quantize=True)
        out[f"{name}/z"] = z.numpy()
        out[f"{name}/image_quantized"] = img_q.numpy()
        out[f"{name}/image_not_quantized"] = img_n.numpy()
        out[f"{name}/indices"] = info[2].numpy().astype(np.int64)
        a, b = O.state_dict_checksum(sd)
        out[f"{name}/sd_checksum"] = np.array([a, b])
    return out


def cond_goldens():
    """Audio conditioning encoder: the REAL ``Conv1DTemporalAttention`` of
    talking_face/ldm/modules/encoders/modules.py (its module-level imports of ``clip`` and ``kornia``, absent here and
    unused by this class, are stubbed) with the synthetic weights of ``oracle.cond_oracle.init_cond_state_dict``."""
    import types
    from oracle import cond_oracle as CO
    R.import_reference("talking_face")
    for name in ("clip", "kornia"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    from ldm.modules.encoders.modules import Conv1DTemporalAttention
    out = {}
    for name, cfg, B, seed in [("tf_audio", CO.TF_AUDIO, 3, 21),
                               ("hidden", dict(seq_len=9, subspace_dim=64, subspace2hidden=True, hidden_dim=48), 2, 22)]:
        sd = CO.init_cond_state_dict(cfg, 0)
        m = Conv1DTemporalAttention(**cfg)
        m.load_state_dict(sd, strict=True)
        m.eval()
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(B, cfg["seq_len"], cfg["subspace_dim"], generator=g)
        with torch.no_grad():
            y = m(x)
        out[f"{name}/x"] = x.numpy()
        out[f"{name}/y"] = y.numpy()
    return out


def vq_encode_goldens():
    """First-stage encode (encode_first_stage of the masked frame) through the REAL reference ``Encoder`` +
    ``quant_conv`` assembled as ``VQModel.__init__`` does (autoencoder.py:36,42), synthetic weights loaded strict=True."""
    import contextlib
    import io
    from oracle import vq_oracle as V
    R.import_reference("face_reenactment")
    from ldm.modules.diffusionmodules.model import Encoder

    class FirstStage(torch.nn.Module):
        def __init__(self, cfg):
            super().__init__()
            dd = cfg["ddconfig"]
            with contextlib.redirect_stdout(io.StringIO()):
                self.encoder = Encoder(**dd)
            self.quant_conv = torch.nn.Conv2d(dd["z_channels"], cfg["embed_dim"], 1)

        def encode(self, x):                                # autoencoder.py:269-272
            return self.quant_conv(self.encoder(x))

    out = {}
    for name, cfg, B, hw, seed in [("tiny_32", V.VQ_TINY, 2, 32, 15), ("tiny_64", V.VQ_TINY, 1, 64, 16),
                                   ("f4_64", V.VQ_F4, 1, 64, 17)]:
        sd = V.init_vq_encoder_state_dict(cfg, 0)
        m = FirstStage(cfg)
        m.load_state_dict(sd, strict=True)
        m.eval()
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(B, 3, hw, hw, generator=g).clamp_(-1, 1)
        with torch.no_grad():
            z = m.encode(x)
        out[f"{name}/x"] = x.numpy()
        out[f"{name}/z"] = z.numpy()
    return out


def c2_trajectory_goldens():
    """BASELINE config 1/2 at full size: the REAL FR UNet, 64x64x3 latents, B=2, the reference ``DDIMSampler.sample``
    (ddim.py:56-162) for 50 steps at eta 0 — final latents and the logged intermediates — plus one B=4 eval at mixed
    timesteps for the batch-invariance check at 64x64 (minutes on CPU)."""
    out = {}
    cfg = O.FR_UNET
    sd = O.init_state_dict(cfg, seed=0)
    ref = R.reference_unet(cfg, "face_reenactment").eval()
    ref.load_state_dict(sd, strict=True)
    model = R.LatentDiffusionStandIn(ref)
    B, hw = 2, 64
    x_T, ctx = synth_inputs(cfg, B, hw, seed=2)
    smp = R.cpu_ddim_sampler(model)
    torch.manual_seed(1234)
    samples, inter = smp.sample(50, batch_size=B, shape=[3, hw, hw], conditioning=ctx, eta=0.0, x_T=x_T,
                                verbose=False, log_every_t=10)
    out["c2/S50_eta0/samples"] = samples.numpy()
    out["c2/S50_eta0/x_inter"] = torch.stack(inter["x_inter"]).numpy()
    out["c2/S50_eta0/pred_x0"] = torch.stack(inter["pred_x0"]).numpy()
    x4, ctx4 = synth_inputs(cfg, 4, hw, seed=3)
    t4 = torch.tensor([981, 501, 261, 21], dtype=torch.long)
    with torch.no_grad():
        out["c2/b4/eps"] = ref(x4, t4, context=ctx4).numpy()
    out["c2/b4/t"] = t4.numpy()
    return out


def c3_chain_goldens():
    """BASELINE config 3: the talking-face progressive sampler exactly as the reference runs it — the private sampler class
    of TF/progressive_sampling_difftalk.py (source-cut, ``oracle.ref_shim.cpu_progressive_sampler``) driving the REAL TF
    UNet, the REAL ``Conv1DTemporalAttention`` (``cond_stage_model_2``) and the REAL first-stage ``Encoder`` + ``quant_conv``
    (``encode_first_stage``; the small VQ_TINY f4 topology keeps the fixture CPU-sized) for 2 autoregressive frames
    (``zid = img``, :316-317): 200 steps at the script's own 32x32 latents, 100 steps at 64x64."""
    import contextlib
    import io
    import types
    from oracle import cond_oracle as CO
    from oracle import vq_oracle as V
    R.import_reference("talking_face")
    for name in ("clip", "kornia"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    from ldm.modules.encoders.modules import Conv1DTemporalAttention
    from ldm.modules.diffusionmodules.model import Encoder

    class FirstStage(torch.nn.Module):
        def __init__(self, cfg):
            super().__init__()
            dd = cfg["ddconfig"]
            with contextlib.redirect_stdout(io.StringIO()):
                self.encoder = Encoder(**dd)
            self.quant_conv = torch.nn.Conv2d(dd["z_channels"], cfg["embed_dim"], 1)

        def encode(self, x):                                # autoencoder.py:269-272
            return self.quant_conv(self.encoder(x))

    out = {}
    cfg = O.TF_UNET
    ref = R.reference_unet(cfg, "talking_face").eval()
    ref.load_state_dict(O.init_state_dict(cfg, seed=0), strict=True)
    aud = Conv1DTemporalAttention(**CO.TF_AUDIO)
    aud.load_state_dict(CO.init_cond_state_dict(CO.TF_AUDIO, 0), strict=True)
    aud.eval()
    fs = FirstStage(V.VQ_TINY)
    fs.load_state_dict(V.init_vq_encoder_state_dict(V.VQ_TINY, 0), strict=True)
    fs.eval()
    model = R.LatentDiffusionStandIn(ref, two_cond=True)
    model.cond_stage_model_2 = aud
    model.encode_first_stage = lambda x: fs.encode(x)
    n_frames, window = 2, 8
    for tag, hw, S in (("hw32_S200", 32, 200), ("hw64_S100", 64, 100)):
        g = torch.Generator().manual_seed(31 + hw)
        c1 = torch.randn(1, 1, 256, generator=g)
        xid = torch.randn(1, 3, hw, hw, generator=g).clamp_(-1, 1)
        xmasks = torch.randn(n_frames, 3, 4 * hw, 4 * hw, generator=g).clamp_(-1, 1)
        audio = torch.randn(n_frames, 768, generator=g)
        draws = []
        smp = R.cpu_progressive_sampler(model, record=draws)
        torch.manual_seed(4321)
        with contextlib.redirect_stdout(io.StringIO()):
            frames, _ = smp.progressive_sampling(c1=c1, xid=xid, xmasks=xmasks, audio_feats=audio, S=S, batch_size=1,
                                                 num_frames=n_frames, shape=[3, hw, hw], audio_window=window, eta=0.,
                                                 verbose=False)
        # the per-step noise_like draws (util.py:264-267) go through the reference's own module; the class body's
        # torch.randn is the per-frame x_T only
        assert len(frames) == n_frames and len(draws) == n_frames, (len(frames), len(draws))
        out[f"c3/{tag}/frames"] = torch.cat(frames).numpy()
        out[f"c3/{tag}/x_T"] = torch.cat(draws).numpy()
    return out


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    only = sys.argv[1:]
    for fn, gen in (("unet.npz", unet_goldens), ("schedule.npz", schedule_goldens),
                    ("trajectory.npz", trajectory_goldens), ("unet64.npz", unet64_goldens),
                    ("manipulation.npz", manipulation_goldens), ("train.npz", train_goldens),
                    ("unet128.npz", unet128_goldens), ("vqdecode.npz", vq_goldens), ("cond.npz", cond_goldens),
                    ("vqencode.npz", vq_encode_goldens),
                    ("c2_trajectory.npz", c2_trajectory_goldens), ("c3_chain.npz", c3_chain_goldens)):
        if only and fn not in only:
            continue
        data = gen()
        path = os.path.join(HERE, fn)
        np.savez_compressed(path, **data)
        print(fn, len(data), "arrays", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
