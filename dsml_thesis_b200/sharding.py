"""Batch-sharded DDIM sampling over the GPUs of one box: one process per GPU, weights replicated, the
samples of a batch split contiguously by rank, NO data-path collective — the UNet has no cross-sample op
(GroupNorm and attention are per sample, openaimodel.py:710-742) and neither has the DDIM update
(ddim.py:165-203) — and ONE collective at the end, the all-gather of the final latents.

The reference samples in a single process on a single GPU (FR/sample_affectnet.py:35,86-120;
TF/progressive_sampling_difftalk.py:350-419), so this is the only multi-GPU seam of the sampling path.  Host logic
only: partitioning, result-invariant noise, the gather.  It is backend-agnostic (NCCL on the box, gloo in the CPU
tests with a stub sampler) and never touches the kernels.
"""
from typing import Dict, Optional, Sequence, Tuple, Union

import torch
import torch.distributed as dist

Cond = Union[None, torch.Tensor, Dict[str, torch.Tensor]]


def _world(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def shard_range(global_batch: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of the global batch owned by ``rank``; the first ``global_batch % world`` ranks own one
    extra sample, so shard sizes differ by at most one and concatenation in rank order restores the batch."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of size {world}")
    base, extra = divmod(int(global_batch), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def initial_noise(seed: int, lo: int, hi: int, shape: Sequence[int], device) -> torch.Tensor:
    """x_T for global samples lo..hi-1.  Sample ``g`` is drawn from its own generator seeded by (seed, g), so the
    latents a sample starts from — and with eta = 0 the whole trajectory — do not depend on how many GPUs share
    the batch (ddim.py:121 draws one ``torch.randn(shape)`` for the whole batch instead)."""
    out = torch.empty((hi - lo, *shape), dtype=torch.float32)
    for i, g in enumerate(range(lo, hi)):
        gen = torch.Generator().manual_seed((int(seed) * 1000003 + g) & 0x7FFFFFFFFFFFFFFF)
        out[i] = torch.randn(tuple(shape), generator=gen)
    return out.to(device)


def shard_conditioning(cond: Cond, lo: int, hi: int, global_batch: int) -> Cond:
    """Rows lo..hi-1 of a conditioning tensor, or of every tensor of a ``ddim2cond`` conditioning dict
    (TF ddim2cond.py:80-81).  A leading dim of 1 is a broadcast conditioning and is kept whole."""
    if cond is None:
        return None
    if isinstance(cond, dict):
        return {k: shard_conditioning(v, lo, hi, global_batch) for k, v in cond.items()}
    if isinstance(cond, (list, tuple)):
        return type(cond)(shard_conditioning(v, lo, hi, global_batch) for v in cond)
    if cond.shape[0] == global_batch:
        return cond[lo:hi]
    if cond.shape[0] == 1:
        return cond
    raise ValueError(f"conditioning has leading dim {cond.shape[0]}, expected {global_batch} (or 1)")


def gather_shards(local: torch.Tensor, global_batch: int, group=None) -> torch.Tensor:
    """The one collective of the sampling path: all-gather of the per-rank results in rank order.  Shards may
    differ in size by one (shard_range), so every rank contributes a buffer padded to the largest shard."""
    rank, world = _world(group)
    if world == 1:
        return local
    sizes = [shard_range(global_batch, world, r) for r in range(world)]
    biggest = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((biggest, *local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    if all(hi - lo == biggest for lo, hi in sizes):
        out = torch.empty((world * biggest, *local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
        return out
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad.contiguous(), group=group)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)], dim=0)


def sample_sharded(sampler, S: int, global_batch: int, shape: Sequence[int], conditioning: Cond = None,
                   seed: int = 0, x_T: Optional[torch.Tensor] = None, gather: bool = True, group=None, **kwargs):
    """``sampler.sample(S, batch_size, shape, conditioning, ...)`` (ddim.py:56-110) for a GLOBAL batch shared by
    all ranks of ``group``.  ``conditioning`` / ``unconditional_conditioning`` / ``x_T`` are GLOBAL tensors (every
    rank passes the same ones, as every rank of a torchrun job runs the same script); each rank samples its
    contiguous shard and — with ``gather`` — receives the full ``[global_batch, C, H, W]`` result.
    Returns ``(samples, per-rank intermediates)`` like ``sample()``."""
    rank, world = _world(group)
    lo, hi = shard_range(global_batch, world, rank)
    device = getattr(getattr(sampler, "model", None), "device", None) or "cpu"
    xt = initial_noise(seed, lo, hi, shape, device) if x_T is None else x_T[lo:hi]
    cond = shard_conditioning(conditioning, lo, hi, global_batch)
    if kwargs.get("unconditional_conditioning") is not None:
        kwargs["unconditional_conditioning"] = shard_conditioning(kwargs["unconditional_conditioning"], lo, hi,
                                                                  global_batch)
    kwargs.setdefault("verbose", False)
    if hi > lo:
        local, inter = sampler.sample(S, batch_size=hi - lo, shape=list(shape), conditioning=cond, x_T=xt, **kwargs)
    else:  # more ranks than samples: this rank only takes part in the gather
        local, inter = torch.empty((0, *shape), dtype=torch.float32, device=xt.device), None
    return (gather_shards(local, global_batch, group) if gather else local), inter
