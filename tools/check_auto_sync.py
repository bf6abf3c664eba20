"""Multi-rank training through the drop-in WITHOUT any explicit gradient sync in user code (what the reference's main.py
does under Lightning DDP, main.py:531-543): every rank starts from DIFFERENT weights and sees DIFFERENT data; after one
loss.backward() the parameters must have been broadcast from rank 0 and every rank must hold the SAME gradients = the
mean over ranks of what each rank computes alone.  Run under torchrun on N GPUs."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import dsml_thesis_b200 as P  # noqa: E402
from dsml_thesis_b200 import train as T  # noqa: E402
from bench import init_weights  # noqa: E402

TINY = dict(image_size=8, in_channels=3, out_channels=3, model_channels=32, attention_resolutions=[1, 2],
            num_res_blocks=1, channel_mult=[1, 2], num_head_channels=32, use_spatial_transformer=True,
            transformer_depth=1, context_dim=64)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)

    def grads_alone(seed_w, seed_x):
        os.environ["LDMB200_AUTO_GRAD_SYNC"] = "1"
        unet = P.UNetModel(**TINY)
        init_weights(unet, seed_w)
        unet = unet.to(dev).train()
        handle = P.DiffusionHandle(unet).to(dev).train()
        g = torch.Generator().manual_seed(seed_x)
        x0, nz = torch.randn(2, 3, 16, 16, generator=g).to(dev), torch.randn(2, 3, 16, 16, generator=g).to(dev)
        ctx, t = torch.randn(2, 1, 64, generator=g).to(dev), torch.tensor([700, 30], device=dev)
        return unet, handle, (x0, ctx, t, nz)

    # (a) the multi-rank step: different initial weights and data per rank, NO GradAllReduce constructed by the user
    unet, handle, (x0, ctx, t, nz) = grads_alone(100 + rank, 200 + rank)
    loss, _ = T.p_losses(handle, x0, ctx, t, nz)
    loss.backward()
    assert unet._grad_sync is not None and unet._grad_sync.auto, "no automatic gradient sync was attached"
    flat = torch.cat([p.grad.reshape(-1) for p in unet.parameters()])
    pflat = torch.cat([p.detach().reshape(-1) for p in unet.parameters()])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    pref = pflat.clone()
    dist.broadcast(pref, src=0)
    assert torch.equal(pflat, pref), "parameters were not broadcast from rank 0"
    assert torch.equal(flat, ref), "gradients differ between ranks after backward"
    # (b) the same step computed rank by rank in ONE process group of size 1 semantics: every rank recomputes each
    # rank's local gradient with rank 0's weights (no sync: a fresh model whose sync is detached), then averages
    acc = torch.zeros_like(flat)
    for r in range(world):
        u2, h2, (x0, ctx, t, nz) = grads_alone(100, 200 + r)
        u2._grad_sync = T.GradAllReduce.__new__(T.GradAllReduce)   # a do-nothing sync: keeps the auto attach away
        u2._grad_sync.auto = False
        u2._grad_bucket_hook = None
        l2, _ = T.p_losses(h2, x0, ctx, t, nz)
        l2.backward()
        acc += torch.cat([p.grad.reshape(-1) for p in u2.parameters()])
    acc /= world
    err = float((flat - acc).norm() / acc.norm())
    if rank == 0:
        print(f"auto grad sync on {world} ranks: parameters broadcast, gradients identical on every rank, "
              f"rel-L2 vs the mean of the per-rank gradients {err:.3e}")
    assert err <= 3e-2, err      # bf16 operand rounding + atomics order: run-to-run ~1e-2
    dist.destroy_process_group()
    if rank == 0:
        print("auto-sync OK")


if __name__ == "__main__":
    main()
