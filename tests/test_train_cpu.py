"""CPU checks of the training path's host logic (no GPU): the two-bucket gradient all-reduce over gloo (world_size 2/3),
the flat-arena bookkeeping and that training on a CPU tensor fails loudly (no fallback)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import dsml_thesis_b200 as P
from dsml_thesis_b200 import train as T
from oracle import ldm_oracle as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, split, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(100 + rank)
        flat = torch.randn(n, generator=g)
        ar = T.GradAllReduce(flat=flat)
        if split is not None:
            ar.bucket_ready(split, n)        # what the backward does once the output half is complete
        ar.finish()
        torch.save(flat, os.path.join(out_dir, f"r{rank}.pt"))
        # a second step without the early bucket must work too (split state is reset)
        flat2 = torch.full((n,), float(rank + 1))
        ar2 = T.GradAllReduce(flat=flat2)
        ar2.finish()
        assert torch.allclose(flat2, torch.full((n,), sum(range(1, world + 1)) / world))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,split", [(2, 700), (2, None), (3, 1), (2, 0)])
def test_two_bucket_gradient_allreduce_is_the_mean_over_ranks(world, split, tmp_path):
    n = 1001
    mp.spawn(_worker, args=(world, _free_port(), n, split, str(tmp_path)), nprocs=world, join=True)
    ref = sum(torch.randn(n, generator=torch.Generator().manual_seed(100 + r)) for r in range(world)) / world
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"r{r}.pt"))
        assert torch.allclose(got, ref, atol=1e-6), r


def _worker_bf16(rank, world, port, n, split, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        flat = torch.randn(n, generator=torch.Generator().manual_seed(200 + rank))
        ar = T.GradAllReduce(flat=flat, wire_dtype=torch.bfloat16)
        if split is not None:
            ar.bucket_ready(split, n)
        ar.finish()
        assert flat.dtype == torch.float32
        torch.save(flat, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,split", [(2, 300), (3, None)])
def test_bf16_wire_gradient_allreduce(world, split, tmp_path):
    """wire_dtype=bfloat16 (SURVEY 8e option): the mean over ranks to bf16 accuracy, identical on every rank, fp32 arena."""
    n = 777
    mp.spawn(_worker_bf16, args=(world, _free_port(), n, split, str(tmp_path)), nprocs=world, join=True)
    ref = sum(torch.randn(n, generator=torch.Generator().manual_seed(200 + r)) for r in range(world)) / world
    got = [torch.load(os.path.join(str(tmp_path), f"r{r}.pt")) for r in range(world)]
    for r in range(world):
        assert torch.equal(got[r], got[0])
        assert float((got[r] - ref).norm() / ref.norm()) <= 1e-2


def test_training_on_cpu_fails_loudly():
    unet = P.UNetModel(**O.TINY_UNET).train()
    x = torch.randn(1, 3, 16, 16)
    with pytest.raises(P.Ldmb200Error):
        unet(x, torch.tensor([5]), context=torch.randn(1, 1, 64))


def test_training_module_never_imports_the_oracle():
    src = open(T.__file__).read()
    assert "oracle" not in src.replace("the CPU oracle", "")
