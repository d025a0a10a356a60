"""Collation of several graphs into one block-diagonal batch, with the layout of torch_geometric's `Batch`
(`DataLoader(dataset, batch_size=B)`, reference graphphysics/train.py:133-165) -- and the
receiver-sorted CSR of the batch assembled from the samples' own CSRs instead of being sorted again.

PyG semantics kept: node-level tensors (`x`, `y`, `pos`, ...) and edge-level tensors (`edge_attr`) are concatenated
along dim 0, `edge_index` along dim 1 with each sample's node offset added, `batch[n]` = sample of node n, `ptr` =
node offsets ([B+1]), `num_graphs` = B.

Why the CSR needs no sort: all nodes of sample g are numbered after those of sample g-1, so the receiver-sorted edge
order of the batch is the concatenation of the samples' receiver-sorted orders.  The topology of a trajectory is
constant, so its CSR is built once (`processor.csr_for` caches it per tensor) and every batch that contains the sample
reuses it: `concat_csr` is four offset-adds and concatenations, `collate(..., csrs=[...])` registers the result for the
batch's `edge_index` so the processor never launches the sort for it.  Pure tensor plumbing: runs on any device.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from .graph import CSRGraph
from .models.processors import Graph

_NODE_KEYS = ("x", "y", "pos")
_EDGE_KEYS = ("edge_attr",)


def concat_csr(csrs: Sequence[CSRGraph]) -> CSRGraph:
    """CSR of the block-diagonal union of graphs given their CSRs, in order (node offsets = running node counts,
    caller edge ids = running edge counts: the edge order of `collate`)."""
    if len(csrs) == 0:
        raise ValueError("concat_csr needs at least one graph")
    dev = csrs[0].rowptr.device
    rowptr, perm, send, recv = [torch.zeros(1, dtype=torch.int32, device=dev)], [], [], []
    n_off = e_off = 0
    for c in csrs:
        if c.rowptr.device != dev:
            raise ValueError("concat_csr: all graphs must be on one device")
        rowptr.append(c.rowptr[1:] + e_off)
        perm.append(c.perm + e_off)
        send.append(c.senders + n_off)
        recv.append(c.receivers + n_off)
        n_off += c.num_nodes
        e_off += c.num_edges
    if n_off >= 2 ** 31 or e_off >= 2 ** 31:
        raise ValueError("concat_csr: the batch exceeds the int32 index range of the kernels")
    return CSRGraph(n_off, e_off, torch.cat(rowptr).to(torch.int32), torch.cat(perm).to(torch.int32),
                    torch.cat(send).to(torch.int32), torch.cat(recv).to(torch.int32),
                    all(c.identity for c in csrs))


def collate(samples: Sequence, csrs: Optional[Sequence[CSRGraph]] = None, extra_node_keys: Sequence[str] = (),
            extra_edge_keys: Sequence[str] = ()) -> Graph:
    """`Batch.from_data_list(samples)` for duck-typed samples (`.x`, `.edge_index`, optional `.y`, `.pos`,
    `.edge_attr`, plus the named extra node- / edge-level tensors).  With `csrs` (one per sample, e.g. from
    `processor.csr_for(sample.edge_index, n)`), the batch's CSR is assembled by `concat_csr`, stored as `batch.csr`
    and registered for `batch.edge_index`."""
    if len(samples) == 0:
        raise ValueError("collate needs at least one sample")
    if csrs is not None and len(csrs) != len(samples):
        raise ValueError("collate: one CSR per sample")
    dev = samples[0].x.device
    counts = [int(s.x.shape[0]) for s in samples]
    ptr = torch.zeros(len(samples) + 1, dtype=torch.int64)
    ptr[1:] = torch.cumsum(torch.tensor(counts, dtype=torch.int64), 0)
    out = {}
    for key in tuple(_NODE_KEYS) + tuple(extra_node_keys) + tuple(_EDGE_KEYS) + tuple(extra_edge_keys):
        vals = [getattr(s, key, None) for s in samples]
        if all(v is None for v in vals):
            continue
        if any(v is None for v in vals):
            raise ValueError(f"collate: `{key}` is missing from some samples")
        out[key] = torch.cat(vals, dim=0)
    out["edge_index"] = torch.cat([s.edge_index + int(ptr[i]) for i, s in enumerate(samples)], dim=1)
    out["batch"] = torch.repeat_interleave(torch.arange(len(samples), device=dev),
                                           torch.tensor(counts, device=dev))
    out["ptr"] = ptr.to(dev)
    out["num_graphs"] = len(samples)
    out["num_nodes"] = int(ptr[-1])
    batch = Graph(**out)
    if csrs is not None:
        for s, c, n in zip(samples, csrs, counts):
            if c.num_nodes != n or c.num_edges != int(s.edge_index.shape[1]):
                raise ValueError("collate: a CSR does not belong to its sample")
        batch.csr = concat_csr(csrs)
        from . import processor as _proc
        _proc.register_csr(batch.edge_index, batch.num_nodes, batch.csr)
    return batch
