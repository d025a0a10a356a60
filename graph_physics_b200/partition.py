"""Node partition of ONE large mesh over several GPUs (BASELINE.json config 5; SURVEY.md 8(e), second row).

The reference is single-device; everything here is new.  Owner-computes by receiver:

* nodes are cut into `world` contiguous ranges with (nearly) equal numbers of in-edges; rank p owns the nodes of
  range p and EVERY edge whose receiver it owns, so edge latents, the segmented aggregation and the node update
  are local and deterministic;
* senders that live on another rank are the rank's halo: read-only copies of their node latents, refreshed once
  per message-passing step (forward), and whose gradient contributions are returned to the owner and summed once
  per step (backward).  Local node numbering: owned nodes first (global id - range begin), then the halo nodes in
  ascending global id (= grouped by owning rank, because ranges are contiguous).

`build_partition` is a set-up function (torch ops on whatever device holds edge_index; called once per topology).
`HaloExchange` moves the rows: one all-to-all-v per direction per step (NCCL through torch.distributed), or -- for
single-process tests -- `LocalHaloExchange`, which serves several partitions of one process with plain copies.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


@dataclass
class Partition:
    rank: int
    world: int
    bounds: torch.Tensor           # [world+1] int64 (CPU): node range of every rank
    num_owned: int
    num_halo: int
    halo_global: torch.Tensor      # [num_halo] int64: global ids of the halo nodes, ascending
    edge_ids: torch.Tensor         # [E_loc] int64: positions of the local edges in the caller's edge_index
    edge_index: torch.Tensor       # [2, E_loc] int64, LOCAL numbering (senders < num_owned + num_halo, receivers < num_owned)
    send_index: torch.Tensor       # [sum(send_counts)] int64: owned local ids to send, grouped by destination rank
    send_counts: List[int]         # rows sent to each rank (forward direction)
    recv_counts: List[int]         # rows received from each rank = halo nodes owned by that rank

    @property
    def node_begin(self) -> int:
        return int(self.bounds[self.rank])

    @property
    def num_ext(self) -> int:
        return self.num_owned + self.num_halo


def balanced_bounds(edge_index: torch.Tensor, num_nodes: int, world: int) -> torch.Tensor:
    """Contiguous node ranges with (nearly) equal in-edge counts: [world+1] int64 on the CPU."""
    deg = torch.bincount(edge_index[1], minlength=num_nodes)
    cum = torch.cumsum(deg, 0)
    total = int(cum[-1]) if num_nodes > 0 else 0
    targets = torch.arange(1, world, device=cum.device, dtype=torch.float64) * (total / world)
    cuts = torch.searchsorted(cum.double(), targets, right=False) + 1
    b = torch.cat([torch.zeros(1, dtype=torch.int64, device=cum.device), cuts.to(torch.int64),
                   torch.full((1,), num_nodes, dtype=torch.int64, device=cum.device)])
    b = torch.cummax(b.clamp_(0, num_nodes), 0).values          # monotone even for degenerate graphs
    return b.cpu()


def _halo_of(edge_index: torch.Tensor, lo: int, hi: int):
    recv = edge_index[1]
    mine = (recv >= lo) & (recv < hi)
    eids = torch.nonzero(mine, as_tuple=False).flatten()
    snd = edge_index[0, eids]
    remote = snd[(snd < lo) | (snd >= hi)]
    return eids, torch.unique(remote)        # unique() sorts ascending


def build_partition(edge_index: torch.Tensor, num_nodes: int, rank: int, world: int,
                    bounds: Optional[torch.Tensor] = None) -> Partition:
    """Partition `rank` of `world`.  Every rank calls this with the same global edge_index (int64 [2,E])."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    ei = edge_index.long()
    if bounds is None:
        bounds = balanced_bounds(ei, num_nodes, world)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    eids, halo = _halo_of(ei, lo, hi)
    n_own, n_halo = hi - lo, int(halo.numel())
    snd, rcv = ei[0, eids], ei[1, eids]
    own_mask = (snd >= lo) & (snd < hi)
    snd_local = torch.where(own_mask, snd - lo, n_own + torch.searchsorted(halo, snd))
    local = torch.stack([snd_local, rcv - lo])
    bdev = bounds.to(halo.device)
    owner = torch.searchsorted(bdev, halo, right=True) - 1        # rank that owns each halo node
    recv_counts = torch.bincount(owner, minlength=world).tolist()
    # what the others need from me: their halo nodes that fall in my range
    send_parts, send_counts = [], []
    for q in range(world):
        if q == rank:
            send_counts.append(0)
            continue
        _, halo_q = _halo_of(ei, int(bounds[q]), int(bounds[q + 1]))
        mineq = halo_q[(halo_q >= lo) & (halo_q < hi)] - lo
        send_parts.append(mineq)
        send_counts.append(int(mineq.numel()))
    send_index = torch.cat(send_parts) if send_parts else torch.zeros(0, dtype=torch.int64, device=ei.device)
    return Partition(rank, world, bounds, n_own, n_halo, halo, eids, local, send_index, send_counts, recv_counts)


class HaloExchange:
    """Rows between ranks over torch.distributed (NCCL on GPUs, gloo on CPU tensors): one all-to-all-v per call."""

    def __init__(self, part: Partition, group=None):
        self.part, self.group = part, group

    def rows_forward(self, send: torch.Tensor, recv: torch.Tensor) -> None:
        """Owners -> halo copies.  send [len(send_index), C]: this rank's owned rows in `send_index` order (grouped by
        destination rank); recv [num_halo, C] (contiguous): receives the halo rows in `halo_global` order."""
        p = self.part
        if p.world == 1:
            return
        dist.all_to_all_single(recv, send, output_split_sizes=p.recv_counts, input_split_sizes=p.send_counts,
                               group=self.group)

    def rows_backward(self, send: torch.Tensor, recv: torch.Tensor) -> None:
        """Halo copies -> owners.  send [num_halo, C]: what this rank holds for its halo rows; recv
        [len(send_index), C]: receives, in `send_index` order, what the other ranks hold for rows this rank owns."""
        p = self.part
        if p.world == 1:
            return
        dist.all_to_all_single(recv, send, output_split_sizes=p.send_counts, input_split_sizes=p.recv_counts,
                               group=self.group)

    def forward(self, x_ext: torch.Tensor) -> None:
        """x_ext [num_ext, C]: fill the halo rows (num_owned:) with the owners' current rows (0:num_owned)."""
        p = self.part
        if p.world == 1:
            return
        self.rows_forward(x_ext[:p.num_owned].index_select(0, p.send_index), x_ext[p.num_owned:])

    def backward(self, g_ext: torch.Tensor) -> None:
        """g_ext [num_ext, C]: add the gradient every rank holds for its halo copies to the owners' rows."""
        p = self.part
        if p.world == 1:
            return
        recv = g_ext.new_empty((int(p.send_index.numel()), g_ext.shape[1]))
        self.rows_backward(g_ext[p.num_owned:].contiguous(), recv)
        g_ext[:p.num_owned].index_add_(0, p.send_index, recv)


class LocalHaloExchange:
    """All partitions live in this process (tests, single-GPU emulation): `forward_all` / `backward_all` take the list of
    every partition's extended array and do what the collectives would."""

    def __init__(self, parts: Sequence[Partition]):
        self.parts = list(parts)

    def forward_all(self, x_exts: Sequence[torch.Tensor]) -> None:
        for p, x in zip(self.parts, x_exts):
            if p.num_halo == 0:
                continue
            owner = torch.searchsorted(p.bounds.to(p.halo_global.device), p.halo_global, right=True) - 1
            for q in range(p.world):
                sel = torch.nonzero(owner == q, as_tuple=False).flatten()
                if sel.numel():
                    src = p.halo_global[sel] - int(p.bounds[q])
                    x[p.num_owned + sel] = x_exts[q][:self.parts[q].num_owned].index_select(0, src)

    def rows_forward_all(self, owned: Sequence[torch.Tensor], halos: Sequence[torch.Tensor]) -> None:
        """owned[q] [num_owned_q, C] (any strides): rows the partitions own; halos[p] [num_halo_p, C] (any strides):
        receive the owners' rows of p's halo nodes."""
        for p, h in zip(self.parts, halos):
            if p.num_halo == 0:
                continue
            owner = torch.searchsorted(p.bounds.to(p.halo_global.device), p.halo_global, right=True) - 1
            for q in range(p.world):
                sel = torch.nonzero(owner == q, as_tuple=False).flatten()
                if sel.numel():
                    h[sel] = owned[q].index_select(0, p.halo_global[sel] - int(p.bounds[q]))

    def rows_backward_all(self, halos: Sequence[torch.Tensor], owned: Sequence[torch.Tensor]) -> None:
        """owned[q] += what every partition p holds in halos[p] for rows q owns (all reads before any add)."""
        adds = []
        for p, h in zip(self.parts, halos):
            if p.num_halo == 0:
                continue
            owner = torch.searchsorted(p.bounds.to(p.halo_global.device), p.halo_global, right=True) - 1
            for q in range(p.world):
                sel = torch.nonzero(owner == q, as_tuple=False).flatten()
                if sel.numel():
                    adds.append((q, p.halo_global[sel] - int(p.bounds[q]), h[sel].clone()))
        for q, idx, val in adds:
            owned[q].index_add_(0, idx, val)

    def backward_all(self, g_exts: Sequence[torch.Tensor]) -> None:
        adds = []
        for p, g in zip(self.parts, g_exts):
            if p.num_halo == 0:
                continue
            owner = torch.searchsorted(p.bounds.to(p.halo_global.device), p.halo_global, right=True) - 1
            for q in range(p.world):
                sel = torch.nonzero(owner == q, as_tuple=False).flatten()
                if sel.numel():
                    adds.append((q, p.halo_global[sel] - int(p.bounds[q]), g[p.num_owned + sel].clone()))
        for q, idx, val in adds:                    # all reads happen before any owner row changes
            g_exts[q][:self.parts[q].num_owned].index_add_(0, idx, val)
