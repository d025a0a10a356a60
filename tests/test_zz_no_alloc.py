"""GPU: `mgn_edge_fwd_hl_noalloc` -- the edge update of steps 1..K-1 with its carry scratch in a file-scope device array
of the library instead of a caller-provided workspace (no allocation inside the call; include/mgn_b200.h) -- called
through the C ABI and compared with the existing entry point `mgn_edge_fwd_hl` on the same inputs, on two calls in a
row (the second on a different, smaller graph: the array keeps the first call's contents, which must not matter).
No counterpart in the reference.  Collected last on purpose (file name),
passes alone in a fresh process and after the other tests.  Tolerance: the one tests/test_gpu_fwd_step.py uses for the
aggregation (2e-5 of the largest entry); the tile images must be identical."""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(golden_dir, copies, seed):
    from graph_physics_b200 import graph as G, ops
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n1 = z["points"].shape[0]
    ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, copies)
    N, E = n1 * copies, ei.shape[1]
    csr = G.build_csr(ei, N)
    dev = "cuda"
    torch.manual_seed(seed)
    ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    b = [torch.randn(128, device=dev) * 0.05 for _ in range(4)]
    sc = 1.0 + 0.1 * torch.randn(128, device=dev)
    packed = ops.pack_block(ew, nw)
    x = torch.randn(N, 128, device=dev)
    e = torch.randn(E, 128, device=dev)
    P = ops.p_fwd(x, packed, b[0])
    hi, lo = ops.e16_image(E, dev), ops.e16_image(E, dev)
    ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e16_next=hi, e16lo_next=lo)
    return dict(N=N, E=E, csr=csr, packed=packed, b=b, sc=sc, P=P, hi=hi, lo=lo)


def _noalloc(s, hi_out, lo_out, e_out, aggr):
    from graph_physics_b200 import _lib as L
    csr, b = s["csr"], s["b"]
    rc = L.lib().mgn_edge_fwd_hl_noalloc(
        L.i64(s["E"]), L.i64(s["N"]), L.p(s["hi"]), L.p(s["lo"]), L.p(hi_out), L.p(lo_out), L.p(e_out), L.p(s["P"]),
        L.p(csr.senders), L.p(csr.receivers), L.p(aggr), L.p(s["packed"]), L.p(b[1]), L.p(b[2]), L.p(b[3]), L.p(s["sc"]),
        L.p(csr.rowptr), L.i32(128), L.stream())
    L.check(rc)


def _compare(s):
    from graph_physics_b200 import ops
    dev, N, E, csr, b = "cuda", s["N"], s["E"], s["csr"], s["b"]
    hi_ref, lo_ref = ops.e16_image(E, dev).zero_(), ops.e16_image(E, dev).zero_()
    e_ref = torch.empty((E, 128), device=dev)
    aggr_ref = ops.edge_fwd_hl(s["hi"], s["lo"], s["P"], csr.senders, csr.receivers, N, s["packed"], b[1], b[2], b[3],
                               s["sc"], hi_out=hi_ref, lo_out=lo_ref, e_out=e_ref, rowptr=csr.rowptr)
    hi_out, lo_out = ops.e16_image(E, dev).zero_(), ops.e16_image(E, dev).zero_()
    e_out = torch.empty((E, 128), device=dev)
    aggr = torch.full((N, 128), float("nan"), device=dev)
    _noalloc(s, hi_out, lo_out, e_out, aggr)
    torch.cuda.synchronize()
    assert torch.isfinite(aggr).all()
    assert torch.equal(hi_out, hi_ref) and torch.equal(lo_out, lo_ref)
    assert float((e_out - e_ref).abs().max()) <= 4e-6 * float(e_ref.abs().max())
    assert float((aggr - aggr_ref).abs().max()) <= 2e-5 * float(aggr_ref.abs().max())
    return bool(torch.equal(aggr, aggr_ref))


def test_no_alloc_entry_point_matches_existing_one_twice(golden_dir):
    from graph_physics_b200 import _lib as L
    L.require_device()
    lib = L.lib()
    lib.mgn_noalloc_max_edges.restype = ctypes.c_int64
    assert int(lib.mgn_noalloc_max_edges()) == 1 << 20
    first = _setup(golden_dir, copies=7, seed=21)       # 77,490 edges, ragged last tile, runs across slot boundaries
    second = _setup(golden_dir, copies=2, seed=22)      # smaller graph: the scratch still holds the first call's rows
    assert first["E"] % 128 != 0 and second["E"] < first["E"]
    same1 = _compare(first)
    same2 = _compare(second)
    same3 = _compare(first)                              # and back again
    print(f"no-alloc edge_fwd_hl vs mgn_edge_fwd_hl: images identical; aggregation bitwise equal: {same1}, {same2}, {same3}")


def test_no_alloc_entry_point_rejects_oversize():
    from graph_physics_b200 import _lib as L
    L.require_device()
    lib = L.lib()
    lib.mgn_noalloc_max_edges.restype = ctypes.c_int64
    too_many = int(lib.mgn_noalloc_max_edges()) + 1
    dummy = torch.zeros(16, device="cuda")
    with pytest.raises(ValueError, match="exceeds the built-in scratch"):
        L.check(lib.mgn_edge_fwd_hl_noalloc(
            L.i64(too_many), L.i64(10), L.p(dummy), L.p(dummy), L.p(None), L.p(None), L.p(None), L.p(dummy), L.p(dummy),
            L.p(dummy), L.p(dummy), L.p(dummy), L.p(dummy), L.p(dummy), L.p(dummy), L.p(dummy), L.p(None), L.i32(128),
            L.stream()))
