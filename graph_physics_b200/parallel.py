"""Multi-GPU plumbing for the processor path (one process per GPU, torch.distributed).

The reference is single-process / single-GPU (train.py:233-235 `devices=1`); everything here is new.

* Batches of meshes (BASELINE.json configs 2-3) are independent units: `shard_meshes` gives each
  rank a contiguous share of the batch, the forward/backward needs NO data-path collective, and the
  replicated weights' gradients are summed with one all-reduce per GraphNetBlock, launched as soon as
  that block's backward is enqueued (`GradSync`, hooked into processor.GRAD_SYNC) so it overlaps the
  remaining blocks' backward over NVLink.
* `allreduce_module_grads` covers parameters outside the fused processor (encoders / decoder).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from . import processor as _proc


def shard_meshes(num_meshes: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[begin, end) of the meshes rank `rank` owns: contiguous, sizes differ by at most one."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    base, rem = divmod(num_meshes, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


class GradSync:
    """Sum (or average) each block's flat gradient over the group, asynchronously.

        sync = GradSync(group=None, average=True); sync.install()
        loss.backward(); sync.wait()        # gradients are reduced when wait() returns
    """

    def __init__(self, group=None, average: bool = True, overlap: bool = True):
        """overlap=True: one asynchronous all-reduce per block as soon as its gradient is complete (call wait()).
        overlap=False: blocking all-reduces enqueued in stream order -- the form a CUDA-graph capture of the step
        can record (work handles cannot be waited on inside a capture); wait() is then a no-op."""
        self.group = group
        self.average = average
        self.overlap = overlap
        self.handles: List = []
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1

    def __call__(self, flat: torch.Tensor, k: int):
        if self.world == 1:
            return
        if self.average:
            flat.mul_(1.0 / self.world)
        if self.overlap:
            self.handles.append(dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
        else:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)

    def install(self):
        _proc.GRAD_SYNC = self
        return self

    def uninstall(self):
        if _proc.GRAD_SYNC is self:
            _proc.GRAD_SYNC = None

    def finish(self):
        """Called by the fused backward before it returns its gradient views: makes the compute stream wait for every
        all-reduce started so far (a stream-side wait, the host does not block), so the values autograd accumulates
        into `.grad` -- also with `zero_grad(set_to_none=False)` or gradient accumulation -- are the reduced ones."""
        for h in self.handles:
            h.wait()
        self.handles.clear()

    def wait(self):
        self.finish()


def allreduce_module_grads(params, group=None, average: bool = True) -> Optional[torch.Tensor]:
    """One flat all-reduce over the .grad of `params` (those that have one); returns the flat buffer."""
    ps = [p for p in params if p.grad is not None]
    if not ps or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return None
    flat = torch.cat([p.grad.reshape(-1) for p in ps])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    if average:
        flat.mul_(1.0 / dist.get_world_size(group))
    off = 0
    for p in ps:
        n = p.grad.numel()
        p.grad.copy_(flat[off:off + n].view_as(p.grad))
        off += n
    return flat


def allreduce_last_grad_blocks(group=None, average: bool = True) -> Optional[torch.Tensor]:
    """One all-reduce over the flat [K, 181504] gradient buffer of the last fused backward sweep
    (processor.LAST_GRAD_BLOCKS; the block parameters' .grad are views of it).  Use this after replaying a
    CUDA-graph capture of the step, where the per-block hook cannot run collectives."""
    buf = _proc.LAST_GRAD_BLOCKS
    if buf is None or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return buf
    dist.all_reduce(buf, op=dist.ReduceOp.AVG if average else dist.ReduceOp.SUM, group=group)
    return buf


def check_grads_alias_last_blocks(params) -> None:
    """`allreduce_last_grad_blocks` reduces the flat buffer of the last backward sweep; that reaches the optimiser only
    if the parameters' `.grad` ARE views of it, i.e. if `.grad` was None before backward (`zero_grad(set_to_none=True)`,
    the PyTorch default).  Raises otherwise (accumulated or pre-existing gradients live in their own storage)."""
    buf = _proc.LAST_GRAD_BLOCKS
    if buf is None:
        raise RuntimeError("no fused backward has run yet")
    lo, hi = buf.data_ptr(), buf.data_ptr() + buf.numel() * buf.element_size()
    bad = [i for i, p in enumerate(params) if p.grad is not None and not (lo <= p.grad.data_ptr() < hi)]
    if bad:
        raise RuntimeError(f"{len(bad)} parameter gradients are not views of processor.LAST_GRAD_BLOCKS: call "
                           "zero_grad(set_to_none=True) before backward, or use GradSync (per-block hook) instead of "
                           "allreduce_last_grad_blocks")
