"""CPU: PyG-`Batch`-layout collation and the CSR of a batch assembled from the samples' CSRs (SURVEY.md 8(f) N2;
reference: `DataLoader(..., batch_size=B)` of torch_geometric, graphphysics/train.py:133-165).  The CSR of the batch
must be bit-identical to the oracle's stable receiver sort of the collated `edge_index`."""
import os

import numpy as np
import pytest
import torch

from graph_physics_b200 import batch as B
from graph_physics_b200 import processor as proc
from graph_physics_b200.graph import CSRGraph
from graph_physics_b200.models import Graph
from oracle import graph_oracle as go


def _csr(ei: torch.Tensor, n: int) -> CSRGraph:
    rowptr, perm, send, recv = go.receiver_csr(ei.numpy(), n)
    ident = bool(np.array_equal(perm, np.arange(perm.shape[0])))
    return CSRGraph(n, int(ei.shape[1]), torch.from_numpy(rowptr), torch.from_numpy(perm), torch.from_numpy(send),
                    torch.from_numpy(recv), ident)


def _samples(golden_dir):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    ei_mesh = torch.as_tensor(z["edge_index"].astype(np.int64))
    n_mesh = int(z["points"].shape[0])
    g = torch.Generator().manual_seed(3)
    out = []
    for n, ei in ((n_mesh, ei_mesh), (7, torch.randint(0, 7, (2, 30), generator=g)), (5, torch.zeros((2, 0), dtype=torch.int64)),
                  (64, torch.randint(0, 64, (2, 129), generator=g)), (n_mesh, ei_mesh)):
        e = ei.shape[1]
        out.append(Graph(x=torch.randn(n, 4, generator=g), y=torch.randn(n, 2, generator=g),
                         pos=torch.randn(n, 2, generator=g), edge_attr=torch.randn(e, 3, generator=g), edge_index=ei))
    return out


def test_collate_has_the_pyg_batch_layout(golden_dir):
    ss = _samples(golden_dir)
    b = B.collate(ss)
    counts = [s.x.shape[0] for s in ss]
    assert b.num_graphs == len(ss) and b.num_nodes == sum(counts)
    assert torch.equal(b.ptr, torch.tensor([0] + list(np.cumsum(counts))))
    assert torch.equal(b.batch, torch.cat([torch.full((c,), i) for i, c in enumerate(counts)]))
    for key in ("x", "y", "pos", "edge_attr"):
        assert torch.equal(getattr(b, key), torch.cat([getattr(s, key) for s in ss]))
    off = 0
    col = 0
    for s, c in zip(ss, counts):
        e = s.edge_index.shape[1]
        assert torch.equal(b.edge_index[:, col:col + e], s.edge_index + off)
        off += c
        col += e
    assert col == b.edge_index.shape[1] and not hasattr(b, "csr")
    with pytest.raises(ValueError):
        B.collate([])
    with pytest.raises(ValueError):
        B.collate(ss, csrs=[_csr(ss[0].edge_index, counts[0])])
    no_y = Graph(x=ss[1].x, edge_index=ss[1].edge_index, edge_attr=ss[1].edge_attr, pos=ss[1].pos)
    with pytest.raises(ValueError):
        B.collate([ss[0], no_y])


def test_batch_csr_from_sample_csrs_is_the_stable_receiver_sort(golden_dir):
    ss = _samples(golden_dir)
    csrs = [_csr(s.edge_index, s.x.shape[0]) for s in ss]
    b = B.collate(ss, csrs=csrs)
    rowptr, perm, send, recv = go.receiver_csr(b.edge_index.numpy(), b.num_nodes)
    c = b.csr
    assert (c.num_nodes, c.num_edges) == (b.num_nodes, b.edge_index.shape[1])
    for got, want in ((c.rowptr, rowptr), (c.perm, perm), (c.senders, send), (c.receivers, recv)):
        assert got.dtype == torch.int32 and np.array_equal(got.numpy(), want)
    assert c.identity is False          # the random samples are not receiver-sorted
    # the mesh fixture IS receiver-sorted?  whatever it is, identity of a batch = all samples' identity
    only_mesh = B.collate([ss[0], ss[4]], csrs=[csrs[0], csrs[4]])
    assert only_mesh.csr.identity == (csrs[0].identity and csrs[4].identity)
    # registered: the processor's lookup returns it without building (building would raise here: no device)
    assert proc.csr_for(b.edge_index, b.num_nodes) is c
    with pytest.raises(ValueError):
        proc.register_csr(b.edge_index, b.num_nodes + 1, c)
    wrong = _csr(ss[1].edge_index, 7)
    with pytest.raises(ValueError):
        B.collate(ss[:2], csrs=[wrong, wrong])


def test_csr_cache_keeps_live_trajectories_over_dead_batches():
    """processor.csr_for / register_csr: entries of tensors that no longer exist go first, then the least recently used;
    a long stream of per-batch registrations must not evict the CSR of a tensor that is still in use."""
    proc._csr_cache.clear()
    keep_ei = torch.randint(0, 9, (2, 20))
    keep = _csr(keep_ei, 9)
    proc.register_csr(keep_ei, 9, keep)
    for _ in range(4 * proc.CSR_CACHE_ENTRIES):                    # short-lived batch tensors
        ei = torch.randint(0, 5, (2, 7))
        proc.register_csr(ei, 5, _csr(ei, 5))
        assert proc.csr_for(keep_ei, 9) is keep
    assert len(proc._csr_cache) <= 2
    live = [torch.randint(0, 5, (2, 7)) for _ in range(proc.CSR_CACHE_ENTRIES + 5)]
    for ei in live:                                                # more live tensors than entries: oldest unused go
        proc.register_csr(ei, 5, _csr(ei, 5))
    assert len(proc._csr_cache) == proc.CSR_CACHE_ENTRIES
    assert id(live[-1]) in proc._csr_cache and id(live[0]) not in proc._csr_cache
    # a tensor modified in place is not a hit (version counter): the lookup falls through to the device build
    proc.register_csr(keep_ei, 9, keep)
    keep_ei[0, 0] = 1
    with pytest.raises(RuntimeError):
        proc.csr_for(keep_ei, 9)
    proc._csr_cache.clear()
