"""Host logic of batch-sharded sampling (dsml_thesis_b200/sharding.py) on CPU: partitioning, result-invariant noise
and the final gather, world_size 2 and 3 over gloo with a stub sampler (the kernels need a B200; the N>1 plumbing
does not)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dsml_thesis_b200 import sharding as SH


class StubSampler:
    """Per-sample independent stand-in for DDIMSampler.sample: out = x_T * 2 + mean(cond) (+ uncond)."""

    class _M:
        device = torch.device("cpu")

    model = _M()

    def sample(self, S, batch_size, shape, conditioning=None, x_T=None, unconditional_conditioning=None, **kw):
        assert x_T.shape == (batch_size, *shape)
        c = conditioning["a"] + conditioning["b"].mean(dim=(1, 2, 3), keepdim=True).view(-1, 1, 1) \
            if isinstance(conditioning, dict) else conditioning
        out = x_T * 2 + c.reshape(c.shape[0], -1).mean(1).view(-1, 1, 1, 1) + S
        if unconditional_conditioning is not None:
            out = out + unconditional_conditioning.reshape(unconditional_conditioning.shape[0], -1).sum(1).view(-1, 1, 1, 1)
        return out, {"x_inter": [x_T]}


def test_shard_range_partitions_the_batch():
    for gb in (0, 1, 5, 64, 257):
        for world in (1, 2, 3, 8):
            spans = [SH.shard_range(gb, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == gb
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        SH.shard_range(4, 2, 2)


def test_initial_noise_is_independent_of_the_partition():
    full = SH.initial_noise(7, 0, 6, (3, 4, 4), "cpu")
    parts = torch.cat([SH.initial_noise(7, lo, hi, (3, 4, 4), "cpu") for lo, hi in ((0, 2), (2, 3), (3, 6))])
    assert torch.equal(full, parts)
    assert not torch.equal(full, SH.initial_noise(8, 0, 6, (3, 4, 4), "cpu"))
    assert abs(float(full.std()) - 1.0) < 0.2


def test_shard_conditioning_forms():
    c = torch.arange(10.0).view(5, 1, 2)
    assert torch.equal(SH.shard_conditioning(c, 1, 3, 5), c[1:3])
    d = SH.shard_conditioning({"x": c, "y": c[:1]}, 3, 5, 5)
    assert torch.equal(d["x"], c[3:5]) and torch.equal(d["y"], c[:1])
    assert SH.shard_conditioning(None, 0, 1, 5) is None
    with pytest.raises(ValueError):
        SH.shard_conditioning(torch.zeros(4, 2), 0, 1, 5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, gb, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(3)
        cond = torch.randn(gb, 1, 8, generator=g)
        uc = torch.randn(gb, 1, 8, generator=g)
        full, inter = SH.sample_sharded(StubSampler(), 4, gb, (3, 4, 4), cond, seed=11, unconditional_conditioning=uc)
        lo, hi = SH.shard_range(gb, world, rank)
        if hi > lo:
            assert inter["x_inter"][0].shape[0] == hi - lo
        dcond = {"a": torch.randn(gb, 1, 8, generator=g), "b": torch.randn(gb, 2, 4, 4, generator=g)}
        full2, _ = SH.sample_sharded(StubSampler(), 2, gb, (3, 4, 4), dcond, seed=5)
        local, _ = SH.sample_sharded(StubSampler(), 2, gb, (3, 4, 4), dcond, seed=5, gather=False)
        assert torch.equal(local, full2[lo:hi])
        torch.save({"full": full, "full2": full2}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,gb", [(2, 6), (2, 5), (3, 7), (3, 2)])
def test_sharded_sampling_matches_single_process(world, gb, tmp_path):
    """Every rank ends with the same full batch, equal to the world_size-1 result (even, ragged and
    fewer-samples-than-ranks splits)."""
    g = torch.Generator().manual_seed(3)
    cond = torch.randn(gb, 1, 8, generator=g)
    uc = torch.randn(gb, 1, 8, generator=g)
    ref, _ = SH.sample_sharded(StubSampler(), 4, gb, (3, 4, 4), cond, seed=11, unconditional_conditioning=uc)
    dcond = {"a": torch.randn(gb, 1, 8, generator=g), "b": torch.randn(gb, 2, 4, 4, generator=g)}
    ref2, _ = SH.sample_sharded(StubSampler(), 2, gb, (3, 4, 4), dcond, seed=5)
    mp.spawn(_worker, args=(world, _free_port(), gb, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"r{r}.pt"))
        assert torch.equal(got["full"], ref), r
        assert torch.equal(got["full2"], ref2), r
