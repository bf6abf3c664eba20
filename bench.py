#!/usr/bin/env python
"""bench.py — DDIM sampling throughput of the B200 denoising path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B] [--hw 64] [--ddim 50]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is ONE full DDIM sampling pass (``--ddim`` UNet evals + fused updates) over one batch of synthetic
latents: the workload of BASELINE config 2 — face_reenactment UNet (in 3, context 512), 64x64x3 latents,
batch 64 per GPU, 50 DDIM steps, eta 0, 16-bit tensor-core mode ("f16": fp16 GEMM operands, bf16 attention
internals, fp32 accumulation — same tensor rate as bf16, meets the <=1e-2 eps bound), random-init weights with every tensor non-zero
(``init`` below; a freshly constructed reference UNet outputs exactly 0).

Prints ONE JSON line (rank 0).  ``value`` = samples/s over all GPUs with inputs resident in HBM; ``e2e`` = the
same through ``DDIMSampler.sample()`` with HOST (pinned) x_T / conditioning and the final latents read back to
the host every step; ``roofline`` = the dominant kernel family (tcgen05 implicit GEMM) timed with CUDA events;
``cpu_baseline`` = the CPU oracle port on this box's host cores over a bounded sample.
``--impl reference`` times that CPU path alone (the reference is pure PyTorch-on-CPU here; /root/reference does
not exist on the GPU box, the oracle restates it and is pinned to it bit-exactly by tests/golden).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FR_UNET = dict(image_size=32, in_channels=3, out_channels=3, model_channels=160, attention_resolutions=[4, 2, 1],
               num_res_blocks=2, channel_mult=[1, 2, 4], num_head_channels=32, use_spatial_transformer=True,
               transformer_depth=1, context_dim=512)
#: algorithmic FLOPs per sample per UNet eval of the REFERENCE op list (probed with hooks)
REF_GFLOP_PER_EVAL = {32: 46.0, 64: 230.0, 128: 1657.0}
METRIC = "DDIM samples/s (64x64x3 latent, 50 steps)"


def metric_name(hw, ddim):
    """BASELINE.json's metric, with the latent size / step count actually run spelled out."""
    return f"DDIM samples/s ({hw}x{hw}x3 latent, {ddim} steps)"


def init_weights(unet, seed=0):
    """Deterministic non-zero synthetic weights (same recipe as oracle.ldm_oracle.init_state_dict, restated here
    because the product path must not import oracle/): U(-1,1)*sqrt(3/fan_in) weights, small biases, norm gains ~1."""
    import torch
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, p in sorted(unet.state_dict().items()):
        shape = tuple(p.shape)
        if len(shape) >= 2:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            sd[name] = ((torch.rand(shape, generator=g) * 2 - 1) / math.sqrt(fan_in) * math.sqrt(3.0)).contiguous()
        elif "norm" in name or name.startswith("out.0") or ".in_layers.0." in name or ".out_layers.0." in name:
            sd[name] = (1.0 + 0.1 * torch.randn(shape, generator=g)) if name.endswith("weight") \
                else 0.1 * torch.randn(shape, generator=g)
        else:
            sd[name] = 0.05 * torch.randn(shape, generator=g)
    unet.load_state_dict(sd, strict=True)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    mx.append(float(f[2]))
                except ValueError:
                    continue
                for n, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def cpu_reference_leg(hw, ddim_steps, steps, warmup, budget_s=25.0):
    """The reference algorithm on the host cores: the oracle port (oracle/ldm_oracle.py, pinned bit-exactly to the
    real reference by tests/golden) at BASELINE config 1's shape — B=2, hw x hw latents.  One 'step' here is ONE
    DDIM step (1 UNet eval + update) of the 50: the bounded sample; samples/s = B / (ddim_steps * s_per_ddim_step)."""
    import torch
    from oracle import ldm_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = dict(FR_UNET)
    sd = O.init_state_dict(cfg, 0)
    g = torch.Generator().manual_seed(0)
    B = 2
    x = torch.randn(B, 3, hw, hw, generator=g)
    ctx = torch.randn(B, 1, 512, generator=g)
    sch = O.ddim_schedule(O.ddpm_buffers()["alphas_cumprod"], ddim_steps, 0.0)
    ts = sch["timesteps"][::-1]

    def one(i, img):
        idx = len(ts) - 1 - i
        t = torch.full((B,), int(ts[i]), dtype=torch.long)
        e = O.unet_forward(sd, cfg, img, t, ctx)
        img, _ = O.ddim_update(img, e, sch["a_t"][idx], sch["a_prev"][idx], sch["sigma_t"][idx],
                               sch["sqrt_one_minus_at"][idx])
        return img

    img = x
    for i in range(warmup):
        img = one(i, img)
    times = []
    t_begin = time.perf_counter()
    for i in range(steps):
        t0 = time.perf_counter()
        img = one(warmup + i, img)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_begin > budget_s and i + 1 >= 1:
            break
    per = sum(times) / len(times)
    value = B / (ddim_steps * per)
    sample = (f"oracle port, B={B}, {hw}x{hw}x3 latent, {len(times)} of {ddim_steps} DDIM steps timed "
              f"({per:.2f} s per step), extrapolated to the full trajectory")
    return value, cores, sample, per, len(times)


TF_UNET = dict(FR_UNET, in_channels=9, context_dim=1024)
TRAIN_METRIC = "DDPM training samples/s (talking_face UNet p_losses fwd+bwd + AdamW/EMA, bf16)"


def cpu_train_reference_leg(hw, budget_steps=1):
    """The reference training arithmetic on the host cores: oracle p_losses (ddpm.py:1014-1047) + autograd through the
    oracle UNet, TF config, B=2 — a bounded sample of BASELINE config 4."""
    import torch
    from oracle import ldm_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = dict(TF_UNET)
    sd = O.init_state_dict(cfg, 0)
    g = torch.Generator().manual_seed(0)
    B = 2
    x0, nz = torch.randn(B, 3, hw, hw, generator=g), torch.randn(B, 3, hw, hw, generator=g)
    ctx, c34 = torch.randn(B, 1, 1024, generator=g), torch.randn(B, 6, hw, hw, generator=g)
    t = torch.tensor([700, 30])
    times = []
    for i in range(budget_steps + 1):
        t0 = time.perf_counter()
        O.p_losses(sd, cfg, x0, t, nz, ctx, c34)
        times.append(time.perf_counter() - t0)
    per = sum(times[1:]) / len(times[1:])
    sample = f"oracle p_losses + autograd (fwd+bwd, no optimizer), TF UNet, B={B}, {hw}x{hw} latents, {len(times) - 1} step(s) after 1 warm-up"
    return B / per, cores, sample, per, len(times) - 1


def train_main(args):
    """BASELINE config 4: talking_face LDM training step (DDPM p_losses fwd/bwd, bf16), batch 32/GPU, NCCL gradient
    all-reduce.  One step = zero_grad, q_sample + UNet forward + loss, backward, (all-reduce), fused AdamW + EMA."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    B, HW = args.batch, args.hw
    workload = (f"C4: talking_face UNet (in 9, ctx 1024, 163M params) p_losses fwd+bwd + AdamW/EMA, {HW}x{HW} latents, "
                f"batch {B}/GPU")
    if args.impl == "reference":
        if rank != 0:
            return
        value, cores, sample, per, n = cpu_train_reference_leg(HW)
        print(json.dumps({"impl": "reference", "metric": TRAIN_METRIC, "value": value, "unit": "samples/s",
                          "n_gpus": args.gpus, "steps": n, "warmup": 1, "ms_per_step": per * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": {"workload": workload, "arm": "reference algorithm on host CPU (oracle port)"},
                          "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
                          "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}))
        return
    import torch
    import torch.distributed as dist
    import dsml_thesis_b200 as P
    from dsml_thesis_b200 import train as T
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    unet = P.UNetModel(**TF_UNET)
    init_weights(unet, 0)
    unet = unet.to(dev).train()
    unet.use_cuda_graph = bool(args.graph)
    handle = P.DiffusionHandle(unet, two_cond=True).to(dev).train()
    opt = T.FusedAdamWEMA(unet, lr=2e-6)          # base_learning_rate of mead-128-ldm-f4.yaml
    wire = {"fp32": None, "bf16": torch.bfloat16}[args.grad_wire]
    sync = T.GradAllReduce(unet, wire_dtype=wire) if world > 1 else None
    g = torch.Generator().manual_seed(2000 + rank)
    host = dict(x0=torch.randn(B, 3, HW, HW, generator=g), nz=torch.randn(B, 3, HW, HW, generator=g),
                c12=torch.randn(B, 1, 1024, generator=g), c34=torch.randn(B, 6, HW, HW, generator=g))
    host = {k: v.pin_memory() for k, v in host.items()}
    t_host = torch.randint(0, 1000, (B,), generator=g).pin_memory()
    devt = {k: v.to(dev) for k, v in host.items()}
    t_dev = t_host.to(dev)
    st = T.train_state(unet)

    def step(d, t):
        opt.zero_grad()
        loss, _ = T.p_losses(handle, d["x0"], d["c12"], t, d["nz"], d["c34"])
        loss.backward()
        if sync is not None:
            sync.finish()
        opt.step()
        return loss

    def step_resident():
        return step(devt, t_dev)

    def step_e2e():
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        loss = step(d, t_host.to(dev, non_blocking=True))
        return float(loss.detach())   # device -> host read of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(3, args.warmup)):
        step_resident()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    st.launches = 0
    ms = timed(step_resident, args.steps)
    n_launch = st.launches + args.steps * 4
    clk = clocks.stop() if rank == 0 else None
    step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    plan = next(iter(st.plans.values()))
    cur = torch.cuda.current_stream()
    st.pack.run(cur.cuda_stream)
    plan.run_timed(cur)
    fwd = plan.run_timed(cur)
    st.flat.zero_()
    plan.run_timed(cur, backward=True)
    st.flat.zero_()
    bwd = plan.run_timed(cur, backward=True)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if os.environ.get("LDMB200_OP_TABLE"):
        with open(os.environ["LDMB200_OP_TABLE"], "w") as f:
            f.write(f"# one training step, B={B}, {HW}x{HW}, bf16: per-launch CUDA-event times\n")
            for phase, ops in (("fwd", fwd), ("bwd", bwd)):
                for i, (name, t_ms, fl, by, what) in enumerate(ops):
                    tf = fl / (t_ms * 1e-3) / 1e12 if t_ms > 0 else 0.0
                    gb = by / (t_ms * 1e-3) / 1e9 if t_ms > 0 else 0.0
                    f.write(f"{phase} {i:4d} {name[8:]:28s} {what:44s} {t_ms:8.4f} ms {tf:8.1f} TF/s {gb:8.0f} GB/s\n")
    fam = {}
    for phase, ops in (("fwd", fwd), ("bwd", bwd)):
        for name, t_ms, fl, by, what in ops:
            key = name + (" dgrad" if what.startswith("dgrad") else "")
            f = fam.setdefault(f"{phase}:{key}", [0.0, 0.0, 0.0, 0])
            f[0] += t_ms
            f[1] += fl
            f[2] += by
            f[3] += 1
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = float(peaks.get("bf16_tflops_sustained", 1400.0))
    dom = max(fam.items(), key=lambda kv: kv[1][0])
    achieved = dom[1][1] / (dom[1][0] * 1e-3) / 1e12 if dom[1][1] else None
    total = args.steps * B * world
    value = total / (ms * 1e-3)
    gflop = {32: 138.0, 64: 690.0}.get(HW)       # 3 x forward: fwd + dgrad + wgrad
    nbytes = sum(v.numel() * 4 for v in host.values()) + t_host.numel() * 8
    line = {
        "metric": TRAIN_METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16 operands, f32 accumulate / residual-stream gradient / optimizer state",
        "data": "synthetic",
        "config": {"workload": workload, "cuda_graph": bool(args.graph),
                   "l2": "per-step working set (saved activations, tens of GB) far exceeds the 126 MB L2",
                   "init": "synthetic non-zero weights, seed 0",
                   "parallelism": f"data-parallel x{world}, flat-arena gradient all-reduce (2 buckets, first overlapped, "
                                  f"{args.grad_wire} on the wire)"},
        "train_tflops_algorithmic": (value * gflop / 1e3) if gflop else None,
        "tensor_frac_of_measured_peak": (value / world * gflop / 1e3 / peak_tf) if gflop else None,
        "e2e": {"value": total / (ms_e2e * 1e-3), "unit": "samples/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 4},
        "gpu_launches": n_launch, "clocks": clk,
        "roofline": {"kernel": dom[0], "bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": (achieved / peak_tf) if achieved else None, "traffic": None,
                     "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback 1.4 PFLOP/s",
                     "launches_per_step": dom[1][3], "ms_per_step": dom[1][0],
                     "how": "the family with the largest share of the step: sum of algorithmic FLOPs / sum of CUDA-event durations"},
        "kernel_breakdown_ms_per_step": {k: round(v[0], 3) for k, v in sorted(fam.items(), key=lambda kv: -kv[1][0])},
        "kernel_tflops": {k: round(v[1] / (v[0] * 1e-3) / 1e12, 1) for k, v in fam.items() if v[1] > 0 and v[0] > 0},
        "step_ms_sum_of_kernels": sum(v[0] for v in fam.values()),
    }
    if not args.no_cpu_baseline:
        v, cores, sample, per, n = cpu_train_reference_leg(min(HW, 32))
        line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


DECODE_METRIC = "first-stage decode images/s (VQ-f4: 64x64x3 latent -> 256x256x3 image)"
VQ_F4 = dict(embed_dim=3, n_embed=16384,
             ddconfig=dict(double_z=False, z_channels=3, resolution=128, in_channels=3, out_ch=3, ch=128,
                           ch_mult=[1, 2, 4], num_res_blocks=2, attn_resolutions=[32], dropout=0.0))


def init_synthetic(module, seed=0, gain=1.0):
    """Non-zero synthetic weights for the first-stage / conditioning modules (sorted-name order, one generator):
    matrices gain*sqrt(3)*U(-1,1)/sqrt(fan_in), N(0,1) codebook, norm gains 1+0.1N, biases 0.05N."""
    import torch
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, prm in sorted(module.named_parameters()):
            if name == "quantize.embedding.weight":
                prm.copy_(torch.randn(prm.shape, generator=g))
            elif prm.dim() >= 2:
                prm.copy_((torch.rand(prm.shape, generator=g) * 2 - 1) * gain * (3.0 / prm[0].numel()) ** 0.5)
            elif "norm" in name and name.endswith("weight"):
                prm.copy_(1.0 + 0.1 * torch.randn(prm.shape, generator=g))
            else:
                prm.copy_(0.05 * torch.randn(prm.shape, generator=g))
    return module


def decode_main(args):
    """--workload decode: VQModelInterface.decode of a batch of latents (nearest-code lookup,
    post_quant_conv, Decoder) - per-frame work of the sampling scripts after the DDIM loop."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    B, HW = args.batch, args.hw
    workload = (f"n2: VQ-f4 first stage (embed 3, 16384 codes, Decoder ch 128 x [1,2,4], attention at the lowest level, "
                f"36.2M params), {HW}x{HW}x3 latents -> {4 * HW}x{4 * HW} images, batch {B}/GPU")
    if args.impl == "reference":
        if rank != 0:
            return
        import torch
        from oracle import vq_oracle as V
        torch.set_num_threads(os.cpu_count() or 1)
        sd = V.init_vq_state_dict(V.VQ_F4, 0)
        hw = min(HW, 32)
        z = torch.randn(1, 3, hw, hw)
        for _ in range(min(1, args.warmup)):
            V.vq_decode(sd, V.VQ_F4, z)
        t0 = time.perf_counter()
        n = max(1, min(args.steps, 3))
        for _ in range(n):
            V.vq_decode(sd, V.VQ_F4, z)
        per = (time.perf_counter() - t0) / n * (HW / hw) ** 2
        value = 1.0 / per
        sample = f"oracle port, B=1, {hw}x{hw} latent, {n} decodes timed, scaled by pixel count to {HW}x{HW}"
        print(json.dumps({"impl": "reference", "metric": DECODE_METRIC, "value": value, "unit": "images/s",
                          "n_gpus": args.gpus, "steps": n, "warmup": args.warmup, "ms_per_step": per * 1e3,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                          "data": "synthetic", "config": {"workload": workload, "arm": "reference algorithm on host CPU (oracle port)"},
                          "cpu_baseline": {"value": value, "unit": "images/s", "cores": os.cpu_count(), "kind": "port", "sample": sample},
                          "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}))
        return
    import torch
    import torch.distributed as dist
    import dsml_thesis_b200 as P
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    m = init_synthetic(P.VQDecoder(**VQ_F4), 0).to(dev).eval()
    m.precision = args.precision
    m.use_cuda_graph = bool(args.graph)
    g = torch.Generator().manual_seed(3000 + rank)
    z_host = torch.randn(B, 3, HW, HW, generator=g).pin_memory()
    z_dev = z_host.to(dev)
    out_host = torch.empty(B, 3, 4 * HW, 4 * HW).pin_memory()

    def step_resident():
        return m.decode(z_dev)

    def step_e2e():
        img = m.decode(z_host.to(dev, non_blocking=True))
        out_host.copy_(img, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(3, args.warmup)):
        step_resident()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ms = timed(step_resident, args.steps)
    clk = clocks.stop() if rank == 0 else None
    step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    plan = next(iter(m._plans.values()))
    cur = torch.cuda.current_stream()
    plan.run_timed(cur)
    ops = plan.run_timed(cur)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if os.environ.get("LDMB200_OP_TABLE"):
        with open(os.environ["LDMB200_OP_TABLE"], "w") as f:
            f.write(f"# one first-stage decode, B={B}, {HW}x{HW} latents, {args.precision}: per-launch CUDA-event times\n")
            for i, (name, t_ms, fl, by, what) in enumerate(ops):
                tf = fl / (t_ms * 1e-3) / 1e12 if t_ms > 0 else 0.0
                gb = by / (t_ms * 1e-3) / 1e9 if t_ms > 0 else 0.0
                f.write(f"{i:4d} {name[8:]:24s} {what:44s} {t_ms:8.4f} ms {tf:8.1f} TF/s {gb:8.0f} GB/s\n")
    fam = {}
    for name, t_ms, fl, by, what in ops:
        f = fam.setdefault(name, [0.0, 0.0, 0.0, 0])
        f[0] += t_ms
        f[1] += fl
        f[2] += by
        f[3] += 1
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = float(peaks.get("bf16_tflops_sustained", 1400.0))
    dom = max(fam.items(), key=lambda kv: kv[1][0])
    achieved = dom[1][1] / (dom[1][0] * 1e-3) / 1e12 if dom[1][1] else None
    flops = sum(v[1] for v in fam.values())
    total = args.steps * B * world
    value = total / (ms * 1e-3)
    line = {
        "metric": DECODE_METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": f"{args.precision} operands, f32 accumulate / residual stream / softmax / code distances",
        "data": "synthetic",
        "config": {"workload": workload, "cuda_graph": bool(args.graph),
                   "l2": "per-decode working set (GBs of activations) far exceeds the 126 MB L2",
                   "init": "synthetic non-zero weights, N(0,1) codebook, seed 0", "parallelism": f"batch-sharded x{world}"},
        "decode_gflop_per_image": flops / B / 1e9,
        "decode_tflops_algorithmic": value / world * flops / B / 1e12,
        "tensor_frac_of_measured_peak": value / world * flops / B / 1e12 / peak_tf,
        "e2e": {"value": total / (ms_e2e * 1e-3), "unit": "images/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": z_host.numel() * 4, "d2h_bytes_per_step": out_host.numel() * 4},
        "gpu_launches": args.steps * (len(ops) + (1 if plan.pack is not None else 0)), "clocks": clk,
        "roofline": {"kernel": dom[0], "bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": (achieved / peak_tf) if achieved else None, "traffic": None,
                     "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback 1.4 PFLOP/s",
                     "launches_per_step": dom[1][3], "ms_per_step": dom[1][0],
                     "how": "the family with the largest share of the decode: sum of algorithmic FLOPs / sum of CUDA-event durations"},
        "kernel_breakdown_ms_per_step": {k: round(v[0], 3) for k, v in sorted(fam.items(), key=lambda kv: -kv[1][0])},
        "step_ms_sum_of_kernels": sum(v[0] for v in fam.values()),
    }
    if not args.no_cpu_baseline:
        from oracle import vq_oracle as V
        torch.set_num_threads(os.cpu_count() or 1)
        sd = V.init_vq_state_dict(V.VQ_F4, 0)
        zc = torch.randn(1, 3, 32, 32)
        V.vq_decode(sd, V.VQ_F4, zc)
        t0 = time.perf_counter()
        V.vq_decode(sd, V.VQ_F4, zc)
        per = (time.perf_counter() - t0) * (HW / 32) ** 2
        line["cpu_baseline"] = {"value": 1.0 / per, "unit": "images/s", "cores": os.cpu_count(), "kind": "port",
                                "sample": f"oracle port, B=1, 32x32 latent, 1 decode timed after 1 warm-up, scaled by pixel "
                                          f"count to {HW}x{HW} (attention scales faster: this favours the CPU)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def torch_eager_gpu_leg(args):
    """--impl torch-eager: the second baseline BASELINE.md names - the reference algorithm as eager PyTorch ops ON THE SAME
    B200 (the oracle's functional restatement, which is bit-identical to the reference modules on CPU; the reference tree
    itself does not travel to the GPU box): fp32 with TF32 off, fp32 with torch's defaults (cuDNN TF32 convolutions), and
    torch.autocast(bfloat16).  Times UNet evals only (the DDIM update is <0.1 % of a step) on config C2."""
    if int(os.environ.get("RANK", "0")) != 0:
        return                                   # a single-GPU baseline: under torchrun only rank 0 runs it
    import torch
    from oracle import ldm_oracle as O
    torch.cuda.set_device(0)
    B, HW, T = args.batch, args.hw, args.ddim
    if args.workload == "train":
        # C4: p_losses forward + autograd backward + torch.optim.AdamW(fused) + a foreach EMA sweep, TF UNet.  No
        # activation checkpointing (the reference always checkpoints its transformer blocks): this favours the baseline.
        g = torch.Generator().manual_seed(2000)
        x0, nz = torch.randn(B, 3, HW, HW, generator=g), torch.randn(B, 3, HW, HW, generator=g)
        ctx, c34 = torch.randn(B, 1, 1024, generator=g).cuda(), torch.randn(B, 6, HW, HW, generator=g).cuda()
        t = torch.randint(0, 1000, (B,), generator=g)
        x_in = torch.cat([O.q_sample(x0, t, nz).cuda(), c34], dim=1)
        nz, t = nz.cuda(), t.cuda()
        out = {}
        for mode in ("fp32_default", "bf16_autocast"):
            torch.backends.cuda.matmul.allow_tf32 = False
            torch.backends.cudnn.allow_tf32 = True
            leaves = {k: v.cuda().requires_grad_(True) for k, v in O.init_state_dict(O.TF_UNET, 0).items()}
            params = list(leaves.values())
            ema = [p.detach().clone() for p in params]
            opt = torch.optim.AdamW(params, lr=2e-6, fused=True)

            def step():
                opt.zero_grad(set_to_none=True)
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "bf16_autocast")):
                    eps = O.unet_forward_autograd(leaves, O.TF_UNET, x_in, t, ctx)
                loss = ((eps.float() - nz) ** 2).mean([1, 2, 3]).mean()
                loss.backward()
                opt.step()
                with torch.no_grad():
                    torch._foreach_lerp_(ema, [p.detach() for p in params], 1.0 - 0.9999)
                return loss
            for _ in range(max(1, args.warmup)):
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(max(1, args.steps)):
                step()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / max(1, args.steps)
            out[mode] = {"step_ms": ms, "samples_per_s": B / (ms * 1e-3)}
            del leaves, params, ema, opt
            torch.cuda.empty_cache()
        best = max(v["samples_per_s"] for v in out.values())
        print(json.dumps({"impl": "torch-eager-gpu", "metric": TRAIN_METRIC, "value": best, "unit": "samples/s", "n_gpus": 1,
                          "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "data": "synthetic",
                          "config": {"workload": f"C4: TF UNet p_losses + autograd + fused AdamW + EMA, {HW}x{HW} latents, "
                                                 f"batch {B} (eager PyTorch ops of the reference algorithm on the same GPU, "
                                                 "no activation checkpointing)"},
                          "modes": out}))
        return
    sd = {k: v.cuda() for k, v in O.init_state_dict(O.FR_UNET, 0).items()}
    g = torch.Generator().manual_seed(1000)
    x = torch.randn(B, 3, HW, HW, generator=g).cuda()
    ctx = torch.randn(B, 1, 512, generator=g).cuda()
    t = torch.full((B,), 501, dtype=torch.long, device="cuda")
    out = {}
    ref = None
    for mode in ("fp32_no_tf32", "fp32_default", "bf16_autocast"):
        torch.backends.cuda.matmul.allow_tf32 = False
        torch.backends.cudnn.allow_tf32 = mode != "fp32_no_tf32"
        ac = torch.autocast("cuda", dtype=torch.bfloat16, enabled=(mode == "bf16_autocast"))
        with torch.no_grad(), ac:
            for _ in range(max(1, args.warmup)):
                y = O.unet_forward(sd, O.FR_UNET, x, t, ctx)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(max(1, args.steps)):
                y = O.unet_forward(sd, O.FR_UNET, x, t, ctx)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / max(1, args.steps)
        if ref is None:
            ref = y.float()
        out[mode] = {"unet_eval_ms": ms, "samples_per_s": B / (T * ms * 1e-3),
                     "eps_rel_l2_vs_fp32": float((y.float() - ref).norm() / ref.norm())}
    best = max(v["samples_per_s"] for v in out.values())
    print(json.dumps({"impl": "torch-eager-gpu", "metric": METRIC, "value": best, "unit": "samples/s", "n_gpus": 1,
                      "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "data": "synthetic",
                      "config": {"workload": f"C2: FR UNet, {HW}x{HW}x3 latents, batch {B}, {T} UNet evals per sample "
                                             "(eager PyTorch ops of the reference algorithm on the same GPU)"},
                      "modes": out}))


C5B_UNET = dict(FR_UNET, channel_mult=[1, 2, 2, 4, 4], attention_resolutions=[4, 8, 16])
C5_GFLOP = {"c5a": 1657.0, "c5b": 375.4}     # algorithmic GFLOP / sample / eval (probed through the reference ctor)


def c5_main(args):
    """--workload c5 (BASELINE config 5, roofline stress): 128x128x3 latents, GLOBAL batch swept 1..512 over the N GPUs
    of the box through sample_sharded (contiguous shards, no data-path collective, final all-gather), both readings of
    SURVEY 8d: c5a = the shipped architecture at 128^2 (self-attention over 16384/4096/1024 tokens), c5b = a 5-level
    channel_mult with attention where the feature map is 32/16/8.  One JSON line PER POINT (rank 0); device time, max
    over ranks; attention and non-attention kernels reported against their own rooflines."""
    import torch
    import torch.distributed as dist
    import dsml_thesis_b200 as P
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    HW, T = 128, args.ddim
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = float(peaks.get("bf16_tflops_sustained", 1400.0))
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    batches = [int(b) for b in args.c5_batches.split(",")]
    max_local = args.c5_max_local

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for reading in args.c5_readings.split(","):
        cfg = FR_UNET if reading == "c5a" else C5B_UNET
        unet = P.UNetModel(**cfg)
        init_weights(unet, 0)
        unet = unet.to(dev).eval()
        unet.precision = args.precision
        unet.use_cuda_graph = bool(args.graph)
        model = P.DiffusionHandle(unet).to(dev)
        smp = P.DDIMSampler(model)
        for GB in batches:
            lo, hi = P.shard_range(GB, world, rank)
            if -(-GB // world) > max_local:
                if rank == 0:
                    print(json.dumps({"workload": reading, "global_batch": GB, "n_gpus": world,
                                      "skipped": f"{-(-GB // world)} samples per GPU exceed --c5-max-local {max_local}"}))
                continue
            gg = torch.Generator().manual_seed(4000 + GB)
            xg = torch.randn(GB, 3, HW, HW, generator=gg).to(dev)
            cg = torch.randn(GB, 1, 512, generator=gg).to(dev)

            def run():
                out, _ = P.sample_sharded(smp, T, GB, [3, HW, HW], conditioning=cg, x_T=xg, eta=0.0)
                return out
            with torch.no_grad():
                clocks = ClockSampler(local_rank)
                for _ in range(max(1, args.warmup)):
                    run()
                barrier()
                if rank == 0:
                    clocks.start()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.steps):
                    run()
                e1.record()
                barrier()
                clk = clocks.stop() if rank == 0 else None
                ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
                if world > 1:
                    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                ms = float(ms.item()) / args.steps
                fam = {}
                plan = unet._plans.get((hi - lo, HW, HW, 1, args.precision)) if hi > lo else None
                if plan is not None:
                    plan.run_timed(torch.cuda.current_stream())
                    for name, t_ms, fl, by, _w in plan.run_timed(torch.cuda.current_stream()):
                        f = fam.setdefault(name, [0.0, 0.0])
                        f[0] += t_ms
                        f[1] += fl
                unet._plans.clear()
            if rank != 0:
                continue
            value = GB / (ms * 1e-3)
            af = fam.get("ldmb200_attention_tc", [0.0, 0.0])
            rest_ms = sum(v[0] for k, v in fam.items() if k != "ldmb200_attention_tc")
            rest_fl = sum(v[1] for k, v in fam.items() if k != "ldmb200_attention_tc")
            mhz = (clk or {}).get("sm_mhz") or 1965.0
            peak_ex2 = 16.0 * sms * mhz * 1e6
            print(json.dumps({
                "metric": metric_name(HW, T), "workload": reading, "value": value, "unit": "samples/s", "n_gpus": world,
                "global_batch": GB, "per_gpu_batch": hi - lo, "steps": args.steps, "warmup": max(1, args.warmup),
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong (fixed global batch split over the ranks)",
                "dtype": args.precision, "data": "synthetic",
                "config": {"workload": f"C5 {reading}: FR UNet " + ("as shipped" if reading == "c5a" else
                                                                     "with channel_mult [1,2,2,4,4], attention_resolutions [4,8,16]")
                           + f", 128x128x3 latents, global batch {GB} over {world} GPU(s), {T}-step DDIM eta=0",
                           "cuda_graph": bool(args.graph)},
                "unet_tflops_algorithmic_per_gpu": value / world * T * C5_GFLOP[reading] / 1e3,
                "tensor_frac_of_measured_peak": value / world * T * C5_GFLOP[reading] / 1e3 / peak_tf,
                "attention": {"ms_per_eval": round(af[0], 4), "tflops": (af[1] / (af[0] * 1e-3) / 1e12) if af[0] else None,
                              "frac_of_ex2_roofline": ((af[1] / 128.0) / (af[0] * 1e-3) / peak_ex2) if af[0] else None},
                "non_attention": {"ms_per_eval": round(rest_ms, 4),
                                  "tflops": (rest_fl / (rest_ms * 1e-3) / 1e12) if rest_ms else None,
                                  "frac_of_tensor_peak": (rest_fl / (rest_ms * 1e-3) / 1e12 / peak_tf) if rest_ms else None},
                "clocks": clk}), flush=True)
        del unet, model, smp
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="sample", choices=["sample", "train", "c3", "decode", "c5"],
                    help="sample = BASELINE config 2 (the headline metric); train = config 4 (DDPM training step); "
                         "c3 = config 3 (talking_face UNet, 200-step progressive DDIM, autoregressive frames)")
    ap.add_argument("--frames", type=int, default=2, help="c3: autoregressive frames per step")
    ap.add_argument("--first-stage", type=int, default=0,
                    help="c3: 1 = also run the per-frame work around the DDIM loop on the GPU path (audio conditioning "
                         "encoder, first-stage encode of the masked frame, first-stage decode of the result)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch-eager"])
    ap.add_argument("--batch", type=int, default=None, help="samples per GPU (weak scaling); default 64 (sample) / 32 (train)")
    ap.add_argument("--hw", type=int, default=64)
    ap.add_argument("--ddim", type=int, default=None, help="DDIM steps: default 50 (sample) / 200 (c3)")
    ap.add_argument("--precision", default="f16", choices=["f16", "bf16", "fp32"])
    ap.add_argument("--graph", type=int, default=int(os.environ.get("LDMB200_CUDA_GRAPH", "1")))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--c5-batches", default="1,2,4,8,16,32,64,128,256,512", help="c5: global batch sizes of the sweep")
    ap.add_argument("--c5-readings", default="c5a,c5b", help="c5: which readings of SURVEY 8d to run")
    ap.add_argument("--c5-max-local", type=int, default=64, help="c5: largest per-GPU batch (activation arena at 128x128)")
    ap.add_argument("--grad-wire", default="fp32", choices=["fp32", "bf16"],
                    help="train, N>1: dtype of the gradient all-reduce on the wire (fp32 = exact, default)")
    args = ap.parse_args()
    if args.batch is None:
        args.batch = 64 if args.workload == "sample" else 16 if args.workload == "decode" else 32
    if args.ddim is None:
        args.ddim = 200 if args.workload == "c3" else 50
    if args.impl == "torch-eager":
        return torch_eager_gpu_leg(args)
    if args.workload == "train":
        return train_main(args)
    if args.workload == "decode":
        return decode_main(args)
    if args.workload == "c5":
        return c5_main(args)
    c3 = args.workload == "c3"

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    workload = (f"C2: face_reenactment UNet (in 3, ctx 512, 156.8M params), {args.hw}x{args.hw}x3 latents, "
                f"batch {args.batch}/GPU, {args.ddim}-step DDIM eta=0")
    metric = metric_name(args.hw, args.ddim)
    if c3:
        workload = (f"C3: talking_face UNet (in 9 = latent + masked-frame + reference-frame latents, audio ctx 1024), "
                    f"{args.hw}x{args.hw} latents, {args.batch} streams/GPU, {args.frames} autoregressive frames x "
                    f"{args.ddim}-step DDIM eta=0 (previous output -> reference channels)")
        metric = f"DDIM frames/s (talking_face, {args.hw}x{args.hw}x3 latent, {args.ddim} steps, progressive)"
        if args.first_stage:
            workload += ("; per frame also Conv1DTemporalAttention over a 17x768 audio window, VQ-f4 encode of the "
                         f"{4 * args.hw}x{4 * args.hw} masked frame and VQ-f4 decode of the generated latent")
            metric = (f"talking-face frames/s ({4 * args.hw}x{4 * args.hw} image in -> image out, {args.ddim}-step DDIM, "
                      "audio conditioning + first-stage encode/decode per frame)")

    if args.impl == "reference":
        if rank != 0:
            return
        value, cores, sample, per, n = cpu_reference_leg(args.hw, args.ddim, max(10, args.steps), max(0, args.warmup),
                                                         budget_s=120.0)
        line = {"impl": "reference", "metric": metric, "value": value, "unit": "samples/s", "n_gpus": args.gpus,
                "steps": n, "warmup": args.warmup, "ms_per_step": per * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": workload, "arm": "reference algorithm on host CPU (oracle port)"},
                "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    import dsml_thesis_b200 as P
    from dsml_thesis_b200 import sampler as S

    assert torch.cuda.is_available(), "bench.py needs a GPU (the product path has no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, HW, T = args.batch, args.hw, args.ddim

    unet = P.UNetModel(**(TF_UNET if c3 else FR_UNET))
    init_weights(unet, 0)
    unet = unet.to(dev).eval()
    unet.precision = args.precision
    unet.use_cuda_graph = bool(args.graph)
    model = P.DiffusionHandle(unet, two_cond=c3).to(dev)
    smp = P.DDIMSampler2Cond(model) if c3 else P.DDIMSampler(model)
    ctx_dim = 1024 if c3 else 512

    # per-rank synthetic inputs (seed 1000 + rank: each rank draws its own batch; throughput does not depend on values)
    g = torch.Generator().manual_seed(1000 + rank)
    x_host = torch.randn(B, 3, HW, HW, generator=g).pin_memory()
    c_host = torch.randn(B, 1, ctx_dim, generator=g).pin_memory()
    c34_host = torch.randn(B, 6, HW, HW, generator=g).pin_memory() if c3 else None
    out_host = torch.empty(B, 3, HW, HW).pin_memory()
    x_dev, c_dev = x_host.to(dev), c_host.to(dev)
    c34_dev = c34_host.to(dev) if c3 else None

    fs = None
    if c3 and args.first_stage:
        aud = init_synthetic(P.Conv1DTemporalAttention(seq_len=17, subspace_dim=768), 1, gain=1.6).to(dev).eval()
        enc = init_synthetic(P.VQEncoder(**VQ_F4), 2).to(dev).eval()
        dec = init_synthetic(P.VQDecoder(**VQ_F4), 0).to(dev).eval()
        enc.precision = dec.precision = args.precision
        model.cond_stage_model_2, model.first_stage_encoder = aud, enc
        fs_host = dict(audio=torch.randn(B, args.frames, 768, generator=g).pin_memory(),
                       masked=torch.randn(args.frames, B, 3, 4 * HW, 4 * HW, generator=g).clamp_(-1, 1).pin_memory())
        fs = {k: v.to(dev) for k, v in fs_host.items()}
        out_host = torch.empty(B, 3, 4 * HW, 4 * HW).pin_memory()

    def progressive(x, c12, c34, fsd=None):
        """progressive_sampling_difftalk.py:286-317: frame f's output becomes the reference-frame latent of f+1.
        With --first-stage the loop IS ``DDIMSampler2Cond.progressive_sampling`` (the method the C3 parity fixture pins against
        the reference's own sampler class): c2 = cond_stage_model_2(audio window), c3 = encode_first_stage(masked frame) per
        frame (286-296), then decode_first_stage of every generated latent (:393)."""
        if fsd is not None:
            frames, _ = smp.progressive_sampling(c1=c12[:, :, :256], xid=c34[:, 3:], xmasks=fsd["masked"],
                                                 audio_feats=fsd["audio"], S=T, batch_size=B, num_frames=args.frames,
                                                 shape=[3, HW, HW], audio_window=8, eta=0., verbose=False,
                                                 x_T=[x] * args.frames)
            img = None
            for lat in frames:
                img = dec.decode(lat)
            return img
        c34 = c34.clone()
        out = None
        for f in range(args.frames):
            cond = {P.DDIMSampler2Cond.CROSS_KEY: c12, P.DDIMSampler2Cond.CONCAT_KEY: c34}
            out, _ = smp.sample(T, batch_size=B, shape=[3, HW, HW], conditioning=cond, eta=0.0, x_T=x, verbose=False)
            c34[:, 3:] = out
        return out

    launches = {"n": 0}
    step_fn = S.ddim_step

    def counted_step(*a, **k):
        launches["n"] += 1
        return step_fn(*a, **k)
    S.ddim_step = counted_step

    def sample_resident():
        if c3:
            out = progressive(x_dev, c_dev, c34_dev, fs)   # streams stay on their rank: no collective
            return out
        out, _ = smp.sample(T, batch_size=B, shape=[3, HW, HW], conditioning=c_dev, eta=0.0, x_T=x_dev, verbose=False)
        if world > 1:
            return P.gather_shards(out, world * B)   # the one collective of the path: final all-gather
        return out

    def sample_e2e():
        x = x_host.to(dev, non_blocking=True)
        c = c_host.to(dev, non_blocking=True)
        if c3:
            out = progressive(x, c, c34_host.to(dev, non_blocking=True),
                              {k: v.to(dev, non_blocking=True) for k, v in fs_host.items()} if fs is not None else None)
        else:
            out, _ = smp.sample(T, batch_size=B, shape=[3, HW, HW], conditioning=c, eta=0.0, x_T=x, verbose=False)
        out_host.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return out_host

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    with torch.no_grad():
        for _ in range(max(3, args.warmup) if not c3 else max(1, args.warmup)):
            sample_resident()
        clocks = ClockSampler(local_rank)
        if rank == 0:
            clocks.start()
        launches["n"] = 0
        unet.launches = 0
        ms = timed(sample_resident, args.steps)
        n_launch = launches["n"] + unet.launches
        clk = clocks.stop() if rank == 0 else None
        for _ in range(2):
            sample_e2e()
        ms_e2e = timed(sample_e2e, args.steps)

        # ---- strong scaling (N > 1): the SAME global batch as one GPU runs, split over the ranks by sample_sharded —
        # the small-per-GPU-batch regime (8 samples per GPU at N = 8) where wave quantisation and per-tile epilogue
        # latency, not a collective, limit scaling.  Global inputs are identical on every rank (seed 999).
        strong = None
        if world > 1 and not c3:
            gg = torch.Generator().manual_seed(999)
            xg = torch.randn(B, 3, HW, HW, generator=gg).to(dev)
            cg = torch.randn(B, 1, ctx_dim, generator=gg).to(dev)

            def sample_strong():
                out, _ = P.sample_sharded(smp, T, B, [3, HW, HW], conditioning=cg, x_T=xg, eta=0.0)
                return out
            for _ in range(2):
                sample_strong()
            ms_strong = timed(sample_strong, args.steps)
            lo, hi = P.shard_range(B, world, rank)
            sp = unet._plans.get((hi - lo, HW, HW, 1, args.precision))
            sfam = {}
            if sp is not None:
                sp.run_timed(torch.cuda.current_stream())
                for name, t_ms, fl, by, _w in sp.run_timed(torch.cuda.current_stream()):
                    f = sfam.setdefault(name, [0.0, 0.0])
                    f[0] += t_ms
                    f[1] += fl
            strong = {"global_batch": B, "per_gpu_batch": hi - lo, "value": args.steps * B / (ms_strong * 1e-3),
                      "unit": "samples/s", "ms_per_step": ms_strong / args.steps,
                      "kernel_breakdown_ms_per_eval": {k: round(v[0], 4) for k, v in sorted(sfam.items(), key=lambda kv: -kv[1][0])},
                      "kernel_tflops": {k: round(v[1] / (v[0] * 1e-3) / 1e12, 1) for k, v in sfam.items() if v[1] > 0 and v[0] > 0},
                      "how": "sample_sharded: contiguous shards of one global batch, no data-path collective, final all_gather; "
                             "device time, max over ranks; breakdown = rank 0's plan at its shard size"}

        # ---- roofline of the dominant kernel family, CUDA events around every launch of one UNet eval
        plan = next(iter(unet._plans.values()))
        if plan.pack is not None:
            plan.pack.run(torch.cuda.current_stream().cuda_stream)
        plan.run_timed(torch.cuda.current_stream())
        per_op = plan.run_timed(torch.cuda.current_stream())
    fam = {}
    folded_flops = 0.0     # algorithmic flops of the folded Upsample convolutions (they execute 4/9 of them)
    for name, t_ms, fl, by, _what in per_op:
        f = fam.setdefault(name, [0.0, 0.0, 0.0, 0])
        f[0] += t_ms
        f[1] += fl
        f[2] += by
        f[3] += 1
        if "upfold" in _what:
            folded_flops += fl
    eval_ms = sum(v[0] for v in fam.values())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = float(peaks.get("bf16_tflops_sustained", 1400.0))
    peak_src = "MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)" if peaks else \
        "fallback 1.4 PFLOP/s sustained (B200_PROFILING.md)"
    dom = "ldmb200_igemm_tc" if args.precision != "fp32" else "ldmb200_igemm_simt"
    # DRAM traffic per kernel family in one eval, from the committed ncu capture of the same workload with the same
    # kernels (profiles/traffic_r02*.json, tools/summarize_traffic.py)
    traffic, traffic_src, traffic_fam = None, None, None
    if B == 64 and HW == 64 and args.precision == "f16" and not c3:
        import glob
        cand = sorted(glob.glob(os.path.join(ROOT, "profiles", "traffic_r*.json")))
        if cand:
            try:
                tj = json.load(open(cand[-1]))
                traffic_src = os.path.relpath(cand[-1], ROOT)
                traffic_fam = {k: v["dram_bytes_total"] for k, v in tj["families"].items()}
                traffic = traffic_fam.get(dom)
            except Exception:
                pass
    d = fam[dom]
    achieved = d[1] / (d[0] * 1e-3) / 1e12
    frames = args.frames if c3 else 1
    total = args.steps * B * world * frames
    value = total / (ms * 1e-3)
    e2e_value = total / (ms_e2e * 1e-3)
    ref_gflop = REF_GFLOP_PER_EVAL.get(HW)
    # second kernel of the step: attention at d_head 32 is bound by the ex2 unit (one exponential per 128 MMA flops,
    # 16 ex2/clk/SM) and the TMEM read port, not by the tensor pipe - report it against THAT roofline
    attn_roof = None
    af = fam.get("ldmb200_attention_tc")
    if af and af[0] > 0:
        mhz = (clk or {}).get("sm_mhz") or (clk or {}).get("sm_max_mhz") or 1965.0
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        peak_ex2 = 16.0 * sms * mhz * 1e6
        ach_ex2 = (af[1] / 128.0) / (af[0] * 1e-3)
        # the same scores are read once from tensor memory as fp32 (4 B each); TMEM read port: 64 B/clk/SM (B300_MICROARCH.md)
        tmem_bytes, peak_tmem = 4.0 * af[1] / 128.0, 64.0 * sms * mhz * 1e6
        attn_roof = {"kernel": "ldmb200_attention_tc", "bound": "tmem-read / mufu (ex2)", "achieved": ach_ex2 / 1e12,
                     "peak": peak_ex2 / 1e12, "unit": "Texp2/s", "frac": ach_ex2 / peak_ex2,
                     "tmem_read": {"achieved": tmem_bytes / (af[0] * 1e-3) / 1e12, "peak": peak_tmem / 1e12, "unit": "TB/s",
                                   "frac": tmem_bytes / (af[0] * 1e-3) / peak_tmem,
                                   "how": "4 B of fp32 score read from tensor memory per exponential (one pass over S); peak = "
                                          "64 B/clk/SM x SM count x the median SM clock under load.  2 of every 8 score pairs take "
                                          "their exponential on the FMA pipe, so the ex2 unit sees 75 % of `achieved`"},
                     "tensor_tflops": af[1] / (af[0] * 1e-3) / 1e12, "ms_per_eval": af[0],
                     "how": "exponentials = attention flops / 128 (4*N^2*32 flops and N^2 exponentials per head); peak = "
                            "16 ex2/clk/SM x SM count x the median SM clock sampled under load"}
    line = {
        "metric": metric, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": {"f16": "f16 operands (bf16 attention internals), f32 accumulate", "bf16": "bf16",
                                      "fp32": "f32"}[args.precision], "data": "synthetic",
        "config": {"workload": workload, "unet_evals_per_step": T, "cuda_graph": bool(args.graph),
                   "l2": "per-eval working set (activation arena + 314 MB 16-bit weights) far exceeds the 126 MB L2; "
                         "no explicit flush",
                   "init": "synthetic non-zero weights, seed 0", "parallelism": f"batch-sharded x{world}, final all_gather"},
        "unet_evals_per_s": value * T / B,
        "unet_tflops_algorithmic": (value * T * ref_gflop / 1e3) if ref_gflop else None,
        "tensor_frac_of_measured_peak": (value / world * T * ref_gflop / 1e3 / peak_tf) if ref_gflop else None,
        "e2e": {"value": e2e_value, "unit": "samples/s", "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": x_host.numel() * 4 + c_host.numel() * 4 + (c34_host.numel() * 4 if c3 else 0)
                + (sum(v.numel() * 4 for v in fs_host.values()) if fs is not None else 0),
                "d2h_bytes_per_step": out_host.numel() * 4},
        "gpu_launches": n_launch + (args.steps * args.frames * (len(next(iter(enc._plans.values())).ops)
                                                                + len(next(iter(dec._plans.values())).ops) + 3)
                                    if fs is not None else 0),
        "clocks": clk,
        "roofline": {"kernel": dom, "bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": achieved / peak_tf, "traffic": traffic, "traffic_source": traffic_src,
                     "executed_achieved": (d[1] - folded_flops * 5.0 / 9.0) / (d[0] * 1e-3) / 1e12,
                     "executed_note": "achieved counts the ALGORITHMIC flops of the reference's op list (SURVEY 8d); the two "
                                      "Upsample convolutions are folded into four 2x2 convolutions that execute 4/9 of theirs: "
                                      "executed_achieved is the rate of the flops the tensor pipe actually ran",
                     "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum summed over the family's launches of "
                                     "one UNet eval (bytes; achieved is likewise the family aggregate per eval)",
                     "peak_source": peak_src,
                     "launches_per_eval": d[3], "ms_per_eval": d[0],
                     "how": "sum of algorithmic FLOPs of the family's launches in one UNet eval / sum of their "
                            "CUDA-event durations on the launching stream"},
        "roofline_attention": attn_roof,
        "traffic_by_family": ({k: {"dram_bytes_per_eval": traffic_fam.get(k), "algorithmic_bytes_per_eval": v[2] or None,
                                   "ms_per_eval": round(v[0], 4)} for k, v in fam.items() if traffic_fam.get(k)}
                              if traffic_fam else None),
        "kernel_breakdown_ms_per_eval": {k: round(v[0], 4) for k, v in sorted(fam.items(), key=lambda kv: -kv[1][0])},
        "eval_ms_sum_of_kernels": eval_ms,
    }
    if strong is not None:
        line["strong"] = strong
    if not args.no_cpu_baseline and not c3:
        v, cores, sample, per, n = cpu_reference_leg(HW, T, 12, 1, budget_s=30.0)
        line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port", "sample": sample}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
