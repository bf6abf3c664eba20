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


# ---- multi-rank training through the reference's own trainer (VERDICT r01 missing 1; main.py:531-543) -----------------
class _StubState:
    def __init__(self, model):
        n = sum(p.numel() for p in model.parameters())
        self.flat = torch.zeros(n)


def _worker_auto_sync(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        unet = P.UNetModel(**O.TINY_UNET).train()
        unet.load_state_dict(O.init_state_dict(O.TINY_UNET, seed=rank), strict=True)   # replicas start DIFFERENT
        stub = _StubState(unet)
        T.train_state = lambda m: stub                       # the real arena needs a GPU; the sync logic does not
        sync = T._ensure_grad_sync(unet)
        assert sync is not None and sync.auto and unet._grad_sync is sync and T._ensure_grad_sync(unet) is sync
        assert unet._grad_bucket_hook == sync.bucket_ready
        # (i) parameters were broadcast from rank 0
        ref = O.init_state_dict(O.TINY_UNET, seed=0)
        for k, v in unet.state_dict().items():
            assert torch.equal(v, ref[k]), k
        # (ii) what the backward node does: bucket 0 while the backward runs, finish() before anyone reads .grad
        stub.flat.fill_(float(rank + 1))
        sync.bucket_ready(stub.flat.numel() // 2, stub.flat.numel())
        sync.finish()
        assert torch.allclose(stub.flat, torch.full_like(stub.flat, sum(range(1, world + 1)) / world))
        # (iii) opting out must raise, not diverge
        unet2 = P.UNetModel(**O.TINY_UNET).train()
        os.environ["LDMB200_AUTO_GRAD_SYNC"] = "0"
        try:
            with pytest.raises(P.Ldmb200Error):
                T._ensure_grad_sync(unet2)
        finally:
            del os.environ["LDMB200_AUTO_GRAD_SYNC"]
        open(os.path.join(out_dir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


def test_multi_rank_training_attaches_a_gradient_sync_automatically(tmp_path):
    mp.spawn(_worker_auto_sync, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")


def test_single_process_training_needs_no_sync():
    unet = P.UNetModel(**O.TINY_UNET).train()
    assert T._ensure_grad_sync(unet) is None


class _LitStandIn(torch.nn.Module):
    """The shape of the reference's LatentDiffusion under Lightning DDP: a UNet inside a wrapper next to other
    trainable parameters (conditioning stage, logvar)."""

    def __init__(self):
        super().__init__()
        self.model = torch.nn.Module()
        self.model.diffusion_model = P.UNetModel(**O.TINY_UNET)
        self.cond_stage_model = torch.nn.Linear(4, 4)

    def forward(self, x):
        return self.cond_stage_model(x).sum()


def _worker_ddp(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        P.install()
        lit = _LitStandIn()
        ddp = torch.nn.parallel.DistributedDataParallel(lit)
        ignored = set(ddp.parameters_to_ignore)
        names = {f"model.diffusion_model.{n}" for n, _ in lit.model.diffusion_model.named_parameters()}
        assert names and names <= ignored
        # DDP manages only the non-UNet parameters: iterating twice must not trip "Expected to have finished reduction"
        for it in range(3):
            ddp.zero_grad()
            ddp(torch.full((2, 4), float(rank + it))).backward()
        g = lit.cond_stage_model.weight.grad.clone()
        dist.all_reduce(g)
        assert torch.allclose(g / world, lit.cond_stage_model.weight.grad)       # averaged by DDP on every rank
        assert all(p.grad is None for p in lit.model.diffusion_model.parameters())
        P.uninstall()
        assert not getattr(torch.nn.parallel.DistributedDataParallel, "_ldmb200_wrapped", False)
        open(os.path.join(out_dir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


def test_install_excludes_the_unet_from_ddp(tmp_path):
    """install() must keep DistributedDataParallel's reducer away from the UNet's parameters (their gradients never pass
    autograd's accumulation hooks), otherwise the second training iteration of main.py's Lightning-DDP run fails — or,
    with find_unused_parameters, silently skips the UNet."""
    mp.spawn(_worker_ddp, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")
