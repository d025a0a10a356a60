"""Host-side multi-rank logic on CPU: world_size-2 gloo process group (no GPU, no kernels).
Covers parallel.shard_meshes, GradSync (the per-block gradient hook the fused backward calls)
and allreduce_module_grads."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from graph_physics_b200 import parallel


def test_shard_meshes_partitions_the_batch():
    for n in (1, 2, 7, 16, 32, 33):
        for w in (1, 2, 3, 4, 8):
            spans = [parallel.shard_meshes(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a0, a1), (b0, b1) in zip(spans[:-1], spans[1:]):
                assert a1 == b0 and a0 <= a1
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        parallel.shard_meshes(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from graph_physics_b200 import processor as proc
        # ---- GradSync: what _ProcessorFn.backward does after each block
        sync = parallel.GradSync(average=True).install()
        assert proc.GRAD_SYNC is sync
        blocks = [torch.full((1000,), float(rank + 1 + 10 * k)) for k in range(3)]
        for k in (2, 1, 0):
            proc.GRAD_SYNC(blocks[k], k)
        sync.wait()
        sync.uninstall()
        assert proc.GRAD_SYNC is None
        ok = all(torch.allclose(blocks[k], torch.full((1000,), 1.5 + 10 * k)) for k in range(3))
        # ---- sum instead of mean
        s2 = parallel.GradSync(average=False)
        t = torch.arange(8, dtype=torch.float32) * (rank + 1)
        s2(t, 0)
        s2.wait()
        ok = ok and torch.equal(t, torch.arange(8, dtype=torch.float32) * 3)
        # ---- parameters outside the fused processor
        lin = torch.nn.Linear(4, 3)
        torch.manual_seed(0)
        for p in lin.parameters():
            p.grad = torch.full_like(p, float(rank))
        frozen = torch.nn.Parameter(torch.zeros(2))       # no grad: skipped
        parallel.allreduce_module_grads(list(lin.parameters()) + [frozen])
        ok = ok and all(torch.allclose(p.grad, torch.full_like(p, 0.5)) for p in lin.parameters())
        ok = ok and frozen.grad is None
        # ---- sharded work adds up: every rank handles its meshes, totals agree after a sum
        b, e = parallel.shard_meshes(7, rank, world)
        cnt = torch.tensor([float(e - b)])
        dist.all_reduce(cnt)
        ok = ok and cnt.item() == 7.0
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_gradsync_world2_gloo():
    world = 2
    port = _free_port()
    mgr = mp.get_context("spawn").Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}
