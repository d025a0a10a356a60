"""Node partition of one mesh over ranks (graph_physics_b200.partition), on CPU: index invariants, the halo exchange
over a world_size-2 gloo group, and -- with the fp64 oracle standing in for the kernels -- that owner-computes with a
per-step halo exchange reproduces the un-partitioned processor, forward and backward."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from graph_physics_b200 import partition as PT
from oracle import mgn_oracle as mo


def _mesh(golden_dir):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    return torch.as_tensor(z["edge_index"].astype(np.int64)), int(z["points"].shape[0])


def _random_graph(n, e, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, n, (2, e), generator=g)


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_partition_invariants(golden_dir, world):
    for ei, n in (_mesh(golden_dir), (_random_graph(300, 2500, 1), 300)):
        parts = [PT.build_partition(ei, n, r, world) for r in range(world)]
        b = parts[0].bounds
        assert int(b[0]) == 0 and int(b[-1]) == n and bool((b[1:] >= b[:-1]).all())
        # every edge belongs to exactly one rank: the owner of its receiver
        all_ids = torch.cat([p.edge_ids for p in parts])
        assert torch.equal(torch.sort(all_ids).values, torch.arange(ei.shape[1]))
        sizes = [int(p.edge_ids.numel()) for p in parts]
        deg_max = int(torch.bincount(ei[1], minlength=n).max())
        assert max(sizes) - min(sizes) <= 2 * deg_max + 1          # balanced by in-edges
        for p in parts:
            lo = p.node_begin
            assert p.num_owned == int(b[p.rank + 1]) - lo and p.num_ext == p.num_owned + p.num_halo
            # local numbering maps back to the global edge list
            to_global = torch.cat([torch.arange(lo, lo + p.num_owned), p.halo_global])
            assert torch.equal(to_global[p.edge_index[0]], ei[0, p.edge_ids])
            assert torch.equal(to_global[p.edge_index[1]], ei[1, p.edge_ids])
            assert bool((p.edge_index[1] < p.num_owned).all())
            # halo = remote senders only, ascending, no duplicates
            assert bool(((p.halo_global < lo) | (p.halo_global >= lo + p.num_owned)).all())
            assert bool((p.halo_global[1:] > p.halo_global[:-1]).all())
            assert sum(p.recv_counts) == p.num_halo and sum(p.send_counts) == int(p.send_index.numel())
            assert p.send_counts[p.rank] == 0 and p.recv_counts[p.rank] == 0
        # what q expects from r is what r sends to q, row for row
        for r in range(world):
            for q in range(world):
                assert parts[r].send_counts[q] == parts[q].recv_counts[r]
                if r == q or parts[r].send_counts[q] == 0:
                    continue
                s0 = sum(parts[r].send_counts[:q])
                sent = parts[r].send_index[s0:s0 + parts[r].send_counts[q]] + parts[r].node_begin
                r0 = sum(parts[q].recv_counts[:r])
                assert torch.equal(sent, parts[q].halo_global[r0:r0 + parts[q].recv_counts[r]])


def test_degenerate_partitions():
    # no edges at all; more ranks than edges; a hub that owns most of the edges
    n = 10
    ei0 = torch.zeros((2, 0), dtype=torch.int64)
    for r in range(2):
        p = PT.build_partition(ei0, n, r, 2)
        assert p.num_halo == 0 and p.edge_ids.numel() == 0
    hub = torch.stack([torch.arange(1, 10), torch.zeros(9, dtype=torch.int64)])
    parts = [PT.build_partition(hub, n, r, 3) for r in range(3)]
    assert sum(p.num_owned for p in parts) == n
    assert sum(int(p.edge_ids.numel()) for p in parts) == 9
    with pytest.raises(ValueError):
        PT.build_partition(hub, n, 3, 3)


def _oracle_partitioned(parts, ei, x, e, p64, steps):
    """The processor with the oracle block per partition and LocalHaloExchange in between, under autograd: the
    exchange is written with differentiable indexing so that backward IS the reverse exchange."""
    xs = [x[p.node_begin:p.node_begin + p.num_owned] for p in parts]
    es = [e[p.edge_ids] for p in parts]
    for k in range(steps):
        full = torch.cat(xs)                                   # contiguous ranges in rank order = global numbering
        nxt = []
        for i, p in enumerate(parts):
            x_ext = torch.cat([xs[i], full[p.halo_global]])
            xo, eo = mo.graph_net_block(x_ext, p.edge_index, es[i], p64, f"processor_list.{k}")
            nxt.append(xo[:p.num_owned])
            es[i] = eo
        xs = nxt
    e_out = torch.empty_like(e).index_copy(0, torch.cat([p.edge_ids for p in parts]), torch.cat(es))
    return torch.cat(xs), e_out


def test_owner_computes_equals_whole_mesh(golden_dir):
    ei, n = _mesh(golden_dir)
    steps, hid = 3, 32
    p = mo.init_params(steps, hidden=hid, seed=2)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(n, hid, generator=g, dtype=torch.float64)
    e = torch.randn(ei.shape[1], hid, generator=g, dtype=torch.float64)
    gx = torch.randn(n, hid, generator=g, dtype=torch.float64)
    res = []
    for world in (1, 3):
        p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}
        xr, er = x.clone().requires_grad_(True), e.clone().requires_grad_(True)
        if world == 1:
            xo, eo = mo.processor(xr, ei, er, p64, steps)
        else:
            parts = [PT.build_partition(ei, n, r, world) for r in range(world)]
            xo, eo = _oracle_partitioned(parts, ei, xr, er, p64, steps)
        (xo * gx).sum().backward()
        res.append((xo.detach(), eo.detach(), xr.grad, er.grad, {k: v.grad for k, v in p64.items()}))
    a, b = res
    assert mo.rel_l2(b[0], a[0]) < 1e-12 and mo.rel_l2(b[1], a[1]) < 1e-12
    assert mo.rel_l2(b[2], a[2]) < 1e-12 and mo.rel_l2(b[3], a[3]) < 1e-12
    assert max(mo.rel_l2(b[4][k], a[4][k]) for k in a[4]) < 1e-10


def test_local_halo_exchange_matches_indexing(golden_dir):
    ei, n = _mesh(golden_dir)
    parts = [PT.build_partition(ei, n, r, 4) for r in range(4)]
    g = torch.Generator().manual_seed(5)
    x = torch.randn(n, 8, generator=g)
    exts = [torch.cat([x[p.node_begin:p.node_begin + p.num_owned], torch.full((p.num_halo, 8), float("nan"))])
            for p in parts]
    hx = PT.LocalHaloExchange(parts)
    hx.forward_all(exts)
    for p, xe in zip(parts, exts):
        assert torch.equal(xe[p.num_owned:], x[p.halo_global])
    # reverse: owners receive the sum of what every other rank holds for them
    gexts = [torch.randn(p.num_ext, 8, generator=g) for p in parts]
    want = torch.cat([ge[:p.num_owned] for p, ge in zip(parts, gexts)]).clone()
    for p, ge in zip(parts, gexts):
        want.index_add_(0, p.halo_global, ge[p.num_owned:])
    hx.backward_all(gexts)
    got = torch.cat([ge[:p.num_owned] for p, ge in zip(parts, gexts)])
    assert torch.allclose(got, want, atol=1e-6)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, golden_dir, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ei, n = _mesh(golden_dir)
        part = PT.build_partition(ei, n, rank, world)
        g = torch.Generator().manual_seed(7)
        x = torch.randn(n, 16, generator=g)
        gfull = [torch.randn(PT.build_partition(ei, n, r, world).num_ext, 16, generator=g) for r in range(world)]
        hx = PT.HaloExchange(part)
        x_ext = torch.cat([x[part.node_begin:part.node_begin + part.num_owned], torch.zeros(part.num_halo, 16)])
        hx.forward(x_ext)
        ok = torch.equal(x_ext[part.num_owned:], x[part.halo_global])
        # reverse exchange against the single-process version on the same data
        parts = [PT.build_partition(ei, n, r, world) for r in range(world)]
        ref = [t.clone() for t in gfull]
        PT.LocalHaloExchange(parts).backward_all(ref)
        mine = gfull[rank].clone()
        hx.backward(mine)
        ok = ok and torch.allclose(mine[:part.num_owned], ref[rank][:part.num_owned], atol=1e-6)
        out.put((rank, bool(ok), part.num_halo))
    finally:
        dist.destroy_process_group()


def test_halo_exchange_gloo_world2(golden_dir):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, golden_dir, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [out.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res) and all(h > 0 for _, _, h in res)


def test_kuhn_grid_edges_equal_face_to_edge_of_its_cells():
    """The config-3/4/5 mesh generator: the edge list built from the 7 Kuhn directions is bit-identical to
    FaceToEdge (oracle restatement) of the tetrahedra, in the same (sender-major) order."""
    from graph_physics_b200 import synth
    from oracle import graph_oracle as go
    for dims in ((2, 2, 2), (3, 3, 3), (4, 5, 6), (7, 3, 2)):
        n = dims[0] * dims[1] * dims[2]
        cells = synth.kuhn_grid_cells(*dims).numpy()
        assert cells.shape == (6 * (dims[0] - 1) * (dims[1] - 1) * (dims[2] - 1), 4)
        want = go.face_to_edge(go.cells_to_faces(cells), n)
        got = synth.kuhn_grid_edges(*dims).numpy()
        assert np.array_equal(got, want)
    assert [synth.grid_for_nodes(w * 159 ** 3 / 8) for w in (1, 2, 4, 8)] == [79, 100, 126, 159]
