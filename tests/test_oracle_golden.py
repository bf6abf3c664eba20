"""Pins the oracle restatement (oracle/ldm_oracle.py) to the committed golden vectors, which are
outputs of the REAL reference (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import ldm_oracle as O
from oracle import ref_shim as R

CASES = {  # name -> (cfg, B, hw, ctx_len, timesteps)
    "tiny_c1": (O.TINY_UNET, 2, 8, 1, [981, 501, 21]),
    "tiny_c3": (O.TINY_UNET, 2, 8, 3, [981, 21]),
    "tiny_hw16": (O.TINY_UNET, 3, 16, 1, [501]),
    "fr_32": (O.FR_UNET, 1, 32, 1, [981]),
    "tf_32": (O.TF_UNET, 1, 32, 1, [501]),
}


def synth_inputs(cfg, B, hw, seed, ctx_len=1):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, cfg["in_channels"], hw, hw, generator=g)
    ctx = torch.randn(B, ctx_len, cfg["context_dim"], generator=g)
    return x, ctx


@pytest.mark.parametrize("name", list(CASES))
def test_unet_eps_matches_reference_golden(name, golden_unet):
    cfg, B, hw, cl, tsteps = CASES[name]
    sd = O.init_state_dict(cfg, 0)
    np.testing.assert_allclose(np.asarray(O.state_dict_checksum(sd)), golden_unet[f"{name}/checksum"], rtol=1e-12)
    assert [len(sd), sum(v.numel() for v in sd.values())] == golden_unet[f"{name}/n_params"].tolist()
    x, ctx = synth_inputs(cfg, B, hw, 1, cl)
    for t in tsteps:
        eps = O.unet_forward(sd, cfg, x, torch.tensor([t] * B), ctx)
        # same ATen ops in the same order: the restatement is bit-exact here; 2e-6 is the fp32
        # thread-count noise floor measured in SURVEY §4 for a differently-threaded host
        assert O.rel_l2(eps, torch.from_numpy(golden_unet[f"{name}/eps_t{t}"])) <= 2e-6
    tm = torch.from_numpy(golden_unet[f"{name}/tmixed"])
    assert O.rel_l2(O.unet_forward(sd, cfg, x, tm, ctx), torch.from_numpy(golden_unet[f"{name}/eps_tmixed"])) <= 2e-6


@pytest.mark.parametrize("name,cfg,t", [("fr_64", O.FR_UNET, 981), ("tf_64", O.TF_UNET, 501)])
def test_unet_eps_64_matches_reference_golden(name, cfg, t, golden_unet64):
    """BASELINE's 64x64 latent size: one eval of the full-size UNets (a few seconds on CPU)."""
    sd = O.init_state_dict(cfg, 0)
    x, ctx = synth_inputs(cfg, 1, 64, 1, 1)
    eps = O.unet_forward(sd, cfg, x, torch.tensor([t]), ctx)
    assert O.rel_l2(eps, torch.from_numpy(golden_unet64[f"{name}/eps_t{t}"])) <= 2e-6


def test_param_inventory_is_626():
    # 626 tensors, 156.76 M (FR) / 163.16 M (TF) parameters
    for cfg, n in ((O.FR_UNET, 156_760_483), (O.TF_UNET, None)):
        shapes = O.param_shapes(cfg)
        assert len(shapes) == 626
        if n:
            assert sum(int(np.prod(s)) for s in shapes.values()) == n


def test_block_taps(golden_unet):
    cfg, B, hw, cl, _ = CASES["tiny_c1"]
    sd = O.init_state_dict(cfg, 0)
    x, ctx = synth_inputs(cfg, B, hw, 1, cl)
    taps = {}
    O.unet_forward(sd, cfg, x, torch.tensor([981] * B), ctx, taps=taps)
    keys = [k for k in golden_unet.files if k.startswith("tiny_c1/tap/")]
    assert len(keys) == len(taps) == 13
    for k in keys:
        assert O.rel_l2(taps[k.split("/tap/")[1]], torch.from_numpy(golden_unet[k])) <= 2e-6


def test_schedules_bit_exact(golden_schedule):
    buf = O.ddpm_buffers()
    for k in ("betas", "alphas_cumprod", "alphas_cumprod_prev"):
        assert np.array_equal(buf[k].numpy(), golden_schedule[f"ddpm/{k}"]), k
    for S in (50, 200, 7):
        for eta in (0.0, 1.0):
            sch = O.ddim_schedule(buf["alphas_cumprod"], S, eta)
            key = f"ddim/S{S}_eta{int(eta)}"
            assert np.array_equal(sch["timesteps"], golden_schedule[key + "/timesteps"])
            for f in ("a_t", "a_prev", "sigma_t", "sqrt_one_minus_at"):
                assert np.array_equal(sch[f], golden_schedule[f"{key}/{f}"]), (S, eta, f)
    # S=50 -> [1,21,...,981]; S=200 -> [1,6,...,996]
    assert O.make_ddim_timesteps(50)[[0, 1, -1]].tolist() == [1, 21, 981]
    assert O.make_ddim_timesteps(200)[[0, 1, -1]].tolist() == [1, 6, 996]


def test_timestep_embedding(golden_schedule):
    t = torch.from_numpy(golden_schedule["temb/t"])
    assert np.array_equal(O.timestep_embedding(t, 160).numpy(), golden_schedule["temb/dim160"])
    assert np.array_equal(O.timestep_embedding(t, 32).numpy(), golden_schedule["temb/dim32"])


@pytest.mark.parametrize("tag,S,eta,scale", [("fr_S10_eta0", 10, 0.0, 1.0), ("fr_S50_eta0", 50, 0.0, 1.0),
                                             ("fr_S10_eta1", 10, 1.0, 1.0), ("fr_S10_cfg3", 10, 0.0, 3.0)])
def test_ddim_trajectory_fr(tag, S, eta, scale, golden_traj):
    cfg = O.TINY_UNET
    sd = O.init_state_dict(cfg, 0)
    x_T = torch.from_numpy(golden_traj["traj/x_T"])
    ctx = torch.from_numpy(golden_traj["traj/ctx"])
    uc = torch.from_numpy(golden_traj["traj/uc"])
    torch.manual_seed(1234)
    noises = [torch.randn(x_T.shape) for _ in range(S)]
    apply_model = lambda x, t, c: O.unet_forward(sd, cfg, x, t, c)
    img, inter, _ = O.ddim_sample(apply_model, O.ddpm_buffers()["alphas_cumprod"], S, x_T, ctx, eta=eta,
                                  unconditional_guidance_scale=scale,
                                  unconditional_conditioning=uc if scale != 1.0 else None,
                                  noises=noises, log_every_t=3)
    assert O.rel_l2(img, torch.from_numpy(golden_traj[f"traj/{tag}/samples"])) <= 1e-5
    assert O.rel_l2(torch.stack(inter["x_inter"]), torch.from_numpy(golden_traj[f"traj/{tag}/x_inter"])) <= 1e-5
    assert O.rel_l2(torch.stack(inter["pred_x0"]), torch.from_numpy(golden_traj[f"traj/{tag}/pred_x0"])) <= 1e-5


def test_ddim_trajectory_tf_two_cond(golden_traj):
    cfg = dict(O.TINY_UNET, in_channels=9)
    sd = O.init_state_dict(cfg, 0)
    x_T = torch.from_numpy(golden_traj["traj/x_T"])
    ctx = torch.from_numpy(golden_traj["traj/ctx"])
    c34 = torch.from_numpy(golden_traj["traj/c34"])
    torch.manual_seed(1234)
    noises = [torch.randn(x_T.shape) for _ in range(10)]
    apply_model = lambda x, t, c: O.unet_forward(sd, cfg, torch.cat([x, c34], 1), t, c)
    img, inter, _ = O.ddim_sample(apply_model, O.ddpm_buffers()["alphas_cumprod"], 10, x_T, ctx, eta=1.0,
                                  noises=noises, log_every_t=3)
    assert O.rel_l2(img, torch.from_numpy(golden_traj["traj/tf_S10_eta1/samples"])) <= 1e-5
    assert O.rel_l2(torch.stack(inter["x_inter"]), torch.from_numpy(golden_traj["traj/tf_S10_eta1/x_inter"])) <= 1e-5


@pytest.mark.skipif(not R.available(), reason="reference tree only exists in the build container")
def test_oracle_bit_exact_vs_live_reference():
    cfg = O.TINY_UNET
    sd = O.init_state_dict(cfg, 0)
    ref = R.reference_unet(cfg).eval()
    ref.load_state_dict(sd, strict=True)
    x, ctx = synth_inputs(cfg, 2, 8, 7, 2)
    t = torch.tensor([3, 777])
    with torch.no_grad():
        assert torch.equal(ref(x, t, context=ctx), O.unet_forward(sd, cfg, x, t, ctx))


# ------------------------------------------------------------------ DDIM inversion / manipulation
@pytest.mark.parametrize("S,strength", [(10, 0.8), (12, 0.6), (50, 0.35)])
def test_strength_schedule_bit_exact(S, strength, golden_manip):
    """ddim2.py make_schedule(strength=) coefficient arrays, incl. sqrt(1 - a_prev) of the inversion step."""
    sch = O.ddim_schedule_strength(O.ddpm_buffers()["alphas_cumprod"], S, strength, 0.0)
    k = f"manip/S{S}_s{int(strength * 100)}"
    assert np.array_equal(sch["timesteps"], golden_manip[k + "/timesteps"])
    for name in ("a_t", "a_prev", "sqrt_one_minus_at", "sqrt_one_minus_a_prev"):
        assert np.array_equal(sch[name], golden_manip[f"{k}/{name}"]), name


def test_strength_one_is_out_of_range_like_the_reference():
    # int(1000 * 1.0) = 1000 indexes past alphas_cumprod in the reference too (ddim2.py:34)
    with pytest.raises(IndexError):
        O.ddim_schedule_strength(O.ddpm_buffers()["alphas_cumprod"], 10, 1.0, 0.0)


@pytest.mark.parametrize("tag,S,strength,scale", [("S10_s80", 10, 0.8, 1.0), ("S12_s60", 12, 0.6, 1.0),
                                                  ("S8_s50_cfg2", 8, 0.5, 2.0)])
def test_latent_manipulation_matches_reference_golden(tag, S, strength, scale, golden_manip):
    """Inversion under c_src + regeneration under c_trg vs the reference's own q_sample_ddim / latent_manipulation."""
    cfg = O.TINY_UNET
    sd = O.init_state_dict(cfg, 0)
    t = lambda k: torch.from_numpy(golden_manip["manip/" + k])
    apply_model = lambda x, tt, c: O.unet_forward(sd, cfg, x, tt, c)
    img, x_inv, _ = O.ddim_latent_manipulation(apply_model, O.ddpm_buffers()["alphas_cumprod"], S, t("x0"), t("c_src"),
                                               t("c_trg"), strength=strength, unconditional_guidance_scale=scale,
                                               unconditional_conditioning=t("uc") if scale != 1.0 else None)
    assert O.rel_l2(x_inv, t(f"{tag}/x_inv")) <= 2e-6
    assert O.rel_l2(img, t(f"{tag}/img")) <= 2e-6


# ---- training step: oracle p_losses + autograd vs the REAL reference UNet's gradients ----
@pytest.fixture(scope="module")
def golden_train():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "train.npz"))


def _train_inputs(cfg, two_cond, B=2, hw=16):
    g = torch.Generator().manual_seed(77)
    x0 = torch.randn(B, 3, hw, hw, generator=g)
    noise = torch.randn(B, 3, hw, hw, generator=g)
    ctx = torch.randn(B, 1, cfg["context_dim"], generator=g)
    c34 = torch.randn(B, 6, hw, hw, generator=g) if two_cond else None
    t = torch.tensor([731, 12][:B])
    return x0, noise, ctx, c34, t


@pytest.mark.parametrize("tag,cfg,two_cond", [("tiny", O.TINY_UNET, False), ("fr16", O.FR_UNET, False),
                                              ("tf16", O.TF_UNET, True)])
def test_p_losses_gradients_match_reference_golden(tag, cfg, two_cond, golden_train):
    x0, noise, ctx, c34, t = _train_inputs(cfg, two_cond)
    assert np.array_equal(O.q_sample(x0, t, noise).numpy(), golden_train[f"train/{tag}/x_noisy"])
    sd = O.init_state_dict(cfg, 0)
    loss, grads = O.p_losses(sd, cfg, x0, t, noise, ctx, c34)
    assert abs(float(loss) - float(golden_train[f"train/{tag}/loss"])) <= 1e-6 * abs(float(loss))
    assert O.rel_l2(grads["context"], torch.from_numpy(golden_train[f"train/{tag}/dctx"])) <= 1e-5
    worst = 0.0
    for name in sd:
        fp = golden_train[f"train/{tag}/gradfp/{name}"]
        gr = grads[name]
        flat = gr.reshape(-1)
        idx = torch.linspace(0, flat.numel() - 1, len(fp) - 1).long()
        mine = torch.cat([gr.double().norm().reshape(1), flat[idx].double()]).numpy()
        scale = max(float(fp[0]), 1e-12)
        worst = max(worst, float(np.abs(mine - fp).max()) / scale)
    assert worst <= 1e-5, worst


def test_unet_eps_128_matches_reference_golden():
    """BASELINE config 5's latent size (128x128: 16384-token self-attention) — one oracle eval vs the real reference."""
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "unet128.npz"))
    sd = O.init_state_dict(O.FR_UNET, 0)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(1, 3, 128, 128, generator=g)
    ctx = torch.randn(1, 1, 512, generator=g)
    eps = O.unet_forward(sd, O.FR_UNET, x, torch.tensor([501]), ctx)
    assert O.rel_l2(eps, torch.from_numpy(gold["fr_128/eps_t501"])) <= 2e-6


# ------------------------------------------------------------------------------------------ first-stage decode (8f n2)
@pytest.mark.parametrize("name,cfg_name", [("tiny_8", "VQ_TINY"), ("tiny_16", "VQ_TINY"), ("f4_16", "VQ_F4")])
def test_vq_decode_matches_reference_golden(name, cfg_name):
    """oracle/vq_oracle.py vs outputs of the real reference VectorQuantizer2 + post_quant_conv + Decoder
    (tests/golden/make_golden.py: vq_goldens): code indices and images bit-exact on CPU."""
    import os
    from conftest import GOLDEN
    from oracle import vq_oracle as V
    g = np.load(os.path.join(GOLDEN, "vqdecode.npz"))
    cfg = getattr(V, cfg_name)
    sd = V.init_vq_state_dict(cfg, 0)
    a, b = O.state_dict_checksum(sd)
    np.testing.assert_allclose([a, b], g[f"{name}/sd_checksum"], rtol=1e-9)
    z = torch.from_numpy(g[f"{name}/z"])
    img, idx = V.vq_decode(sd, cfg, z)
    np.testing.assert_array_equal(idx.numpy(), g[f"{name}/indices"])
    assert O.rel_l2(img, torch.from_numpy(g[f"{name}/image_quantized"])) <= 1e-6
    img, none = V.vq_decode(sd, cfg, z, force_not_quantize=True)
    assert none is None
    assert O.rel_l2(img, torch.from_numpy(g[f"{name}/image_not_quantized"])) <= 1e-6


# ------------------------------------------------------------------------------------------ audio conditioning (8f n4)
@pytest.mark.parametrize("name,cfg", [("tf_audio", None), ("hidden", dict(seq_len=9, subspace_dim=64, subspace2hidden=True,
                                                                          hidden_dim=48))])
def test_conv1d_temporal_attention_matches_reference_golden(name, cfg):
    """oracle/cond_oracle.py vs the output of the real reference Conv1DTemporalAttention (make_golden.py: cond_goldens)."""
    import os
    from conftest import GOLDEN
    from oracle import cond_oracle as CO
    g = np.load(os.path.join(GOLDEN, "cond.npz"))
    cfg = cfg or CO.TF_AUDIO
    sd = CO.init_cond_state_dict(cfg, 0)
    y = CO.conv1d_temporal_attention(sd, cfg, torch.from_numpy(g[f"{name}/x"]))
    assert y.shape == g[f"{name}/y"].shape
    assert O.rel_l2(y, torch.from_numpy(g[f"{name}/y"])) <= 1e-6


@pytest.mark.parametrize("name,cfg_name", [("tiny_32", "VQ_TINY"), ("tiny_64", "VQ_TINY"), ("f4_64", "VQ_F4")])
def test_vq_encode_matches_reference_golden(name, cfg_name):
    """oracle/vq_oracle.py: vq_encode vs the real reference Encoder + quant_conv (make_golden.py: vq_encode_goldens)."""
    import os
    from conftest import GOLDEN
    from oracle import vq_oracle as V
    g = np.load(os.path.join(GOLDEN, "vqencode.npz"))
    cfg = getattr(V, cfg_name)
    z = V.vq_encode(V.init_vq_encoder_state_dict(cfg, 0), cfg, torch.from_numpy(g[f"{name}/x"]))
    assert O.rel_l2(z, torch.from_numpy(g[f"{name}/z"])) <= 1e-6
