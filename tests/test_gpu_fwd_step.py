"""GPU: one message-passing step of the fused forward kernels vs the fp32/fp64 oracle.
Tolerance (BASELINE.json north_star): relative L2 <= 1e-3 per processor step."""
import os

import numpy as np
import pytest
import torch

from oracle import graph_oracle as go
from oracle import mgn_oracle as mo
from oracle.synth import processor_inputs

pytestmark = pytest.mark.gpu
TOL_STEP = 1e-3


def _block_params(seed, dtype=torch.float32):
    return mo.init_params(1, hidden=128, seed=seed, dtype=dtype)


def _run_step(ei_np, n, x0, e0, p, use_rowmap):
    from graph_physics_b200 import graph as G, ops
    dev = "cuda"
    ei = torch.as_tensor(ei_np, dtype=torch.int64, device=dev)
    csr = G.build_csr(ei, n)
    pre = "processor_list.0."
    g = {k: v.to(dev) for k, v in p.items()}
    ew = [g[pre + f"edge_block.{i}.weight"] for i in (0, 2, 4, 6)]
    nw = [g[pre + f"node_block.{i}.weight"] for i in (0, 2, 4, 6)]
    packed = ops.pack_block(ew, nw)
    xd, ed = x0.to(dev), e0.to(dev)
    P = ops.p_fwd(xd, packed, g[pre + "edge_block.0.bias"])
    if use_rowmap:
        e_in, rowmap = ed, csr.perm
    else:
        e_in, rowmap = ed.index_select(0, csr.perm.long()).contiguous(), None
    e_out, aggr = ops.edge_fwd(e_in, rowmap, P, csr.senders, csr.receivers, n, packed,
                               g[pre + "edge_block.2.bias"], g[pre + "edge_block.4.bias"],
                               g[pre + "edge_block.6.bias"], g[pre + "edge_block.7.scale"])
    x_out = ops.node_fwd(xd, aggr, packed, g[pre + "node_block.0.bias"], g[pre + "node_block.2.bias"],
                         g[pre + "node_block.4.bias"], g[pre + "node_block.6.bias"], g[pre + "node_block.7.scale"])
    torch.cuda.synchronize()
    if not use_rowmap:
        back = torch.empty_like(e_out)
        back[csr.perm.long()] = e_out
        e_out = back
    return x_out.cpu(), e_out.cpu(), aggr.cpu(), P.cpu(), csr


@pytest.mark.parametrize("use_rowmap", [True, False])
def test_cylinder_one_step(golden_dir, use_rowmap):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    ei = z["edge_index"].astype(np.int64)
    n, e = 1923, ei.shape[1]
    p = _block_params(0)
    x0, e0, _ = processor_inputs(n, e, 128, seed=0)
    x1, e1, aggr, P, _ = _run_step(ei, n, x0, e0, p, use_rowmap)
    p64 = {k: v.double() for k, v in p.items()}
    xr, er = mo.graph_net_block(x0.double(), torch.tensor(ei), e0.double(), p64, "processor_list.0")
    # intermediate checks make failures easy to localise
    W1 = p64["processor_list.0.edge_block.0.weight"]
    Pref = torch.cat([x0.double() @ W1[:, 128:256].t() + p64["processor_list.0.edge_block.0.bias"],
                      x0.double() @ W1[:, 256:384].t()], dim=1)
    assert mo.rel_l2(P.double(), Pref) < TOL_STEP
    e_ = er - e0.double()
    aggr_ref = torch.zeros(n, 128, dtype=torch.float64).index_add_(0, torch.tensor(ei[1]), e_)
    assert mo.rel_l2(e1.double() - e0.double(), e_) < 2e-3      # the MLP output alone
    assert mo.rel_l2(aggr.double(), aggr_ref) < 2e-3
    assert mo.rel_l2(e1, er) < TOL_STEP
    assert mo.rel_l2(x1, xr) < TOL_STEP


@pytest.mark.parametrize("n,e", [(4, 8), (5, 1), (130, 129), (300, 1000), (1000, 257)])
def test_ragged_random_graphs(n, e):
    g = torch.Generator().manual_seed(n * 1000 + e)
    ei = torch.randint(0, n, (2, e), generator=g).numpy()
    p = _block_params(1)
    x0, e0, _ = processor_inputs(n, e, 128, seed=2)
    x1, e1, aggr, P, _ = _run_step(ei, n, x0, e0, p, True)
    p64 = {k: v.double() for k, v in p.items()}
    xr, er = mo.graph_net_block(x0.double(), torch.tensor(ei), e0.double(), p64, "processor_list.0")
    assert mo.rel_l2(e1, er) < TOL_STEP
    assert mo.rel_l2(x1, xr) < TOL_STEP


def test_no_edges():
    p = _block_params(1)
    x0, e0, _ = processor_inputs(10, 0, 128, seed=2)
    x1, e1, aggr, P, _ = _run_step(np.zeros((2, 0), np.int64), 10, x0, e0, p, True)
    p64 = {k: v.double() for k, v in p.items()}
    xr, er = mo.graph_net_block(x0.double(), torch.zeros((2, 0), dtype=torch.int64), e0.double(), p64,
                                "processor_list.0")
    assert e1.shape[0] == 0 and float(aggr.abs().max()) == 0.0
    assert mo.rel_l2(x1, xr) < TOL_STEP


def _hl_setup(golden_dir, copies, seed):
    import numpy as np
    from graph_physics_b200 import graph as G, ops
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n1 = z["points"].shape[0]
    ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, copies)
    N, E = n1 * copies, ei.shape[1]
    csr = G.build_csr(ei, N)
    torch.manual_seed(seed)
    dev = "cuda"
    ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    b = [torch.randn(128, device=dev) * 0.05 for _ in range(4)]
    sc = 1.0 + 0.1 * torch.randn(128, device=dev)
    packed = ops.pack_block(ew, nw)
    x = torch.randn(N, 128, device=dev)
    e = torch.randn(E, 128, device=dev)
    P = ops.p_fwd(x, packed, b[0])
    return ops, csr, N, E, packed, b, sc, P, e


def _image_rows(img, E):
    from graph_physics_b200 import ops
    return ops.image_rows(img, E)


@pytest.mark.parametrize("copies", [1, 7])
def test_image_remainder_edge_kernel_matches_row_kernel(golden_dir, copies):
    """csrc/edge_fwd_hl.cu (edge stream as fp16 image + fp16 remainder, bulk copies only, row-owner output and shuffle
    segmented sum) against mlp_fwd_kernel<true> (fp32 rows) on the same step: the products see identical operands, so
    the MLP output is identical; the residual add sees e to >= 21 bits, so rows agree to ~1e-6 and the images of the
    output describe the same rows; the aggregation differs only in summation order."""
    ops, csr, N, E, packed, b, sc, P, e = _hl_setup(golden_dir, copies, 4)
    dev = "cuda"
    hi, lo = ops.e16_image(E, dev), ops.e16_image(E, dev)
    # step A (row kernel) produces e1 as rows and as image + remainder
    e1, _ = ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e16_next=hi,
                         e16lo_next=lo)
    assert float((_image_rows(hi, E) + _image_rows(lo, E) - e1).abs().max()) <= 2.0 ** -20 * float(e1.abs().max())
    # the same, rows not written at all
    hi_b, lo_b = ops.e16_image(E, dev), ops.e16_image(E, dev)
    none_rows, _ = ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc,
                                e16_next=hi_b, e16lo_next=lo_b, write_rows=False)
    assert none_rows is None and torch.equal(hi_b, hi) and torch.equal(lo_b[: E // 128], lo[: E // 128])
    # step B on e1: row kernel (image-fed) vs image + remainder kernel
    hi2_ref = ops.e16_image(E, dev)
    e2_ref, aggr_ref = ops.edge_fwd(e1, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc,
                                    e16_in=hi, e16_next=hi2_ref)
    hi2, lo2 = ops.e16_image(E, dev), ops.e16_image(E, dev)
    e2 = torch.empty_like(e1)
    aggr = ops.edge_fwd_hl(hi, lo, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, hi_out=hi2,
                           lo_out=lo2, e_out=e2)
    torch.cuda.synchronize()
    scale_e = float(e2_ref.abs().max())
    assert float((e2 - e2_ref).abs().max()) <= 4e-6 * scale_e
    assert float((aggr - aggr_ref).abs().max()) <= 2e-5 * float(aggr_ref.abs().max())
    r_hi, r_lo = _image_rows(hi2, E), _image_rows(lo2, E)
    assert float((r_hi + r_lo - e2).abs().max()) <= 2.0 ** -20 * scale_e
    assert float((r_hi - e2).abs().max()) <= 2.0 ** -10 * scale_e          # hi alone is fp16(e2)
    # rows past the end of the last tile are zeros in both images
    if E % 128:
        tail_hi = _image_rows(hi2.clone(), (E + 127) // 128 * 128)[E:]
        tail_lo = _image_rows(lo2.clone(), (E + 127) // 128 * 128)[E:]
        assert float(tail_hi.abs().max()) == 0.0 and float(tail_lo.abs().max()) == 0.0
    # images only (what steps 1..K-2 of the processor loop ask for), and aggregation only (the last step)
    hi3, lo3 = ops.e16_image(E, dev), ops.e16_image(E, dev)
    aggr3 = ops.edge_fwd_hl(hi, lo, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, hi_out=hi3,
                            lo_out=lo3)
    aggr4 = ops.edge_fwd_hl(hi, lo, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc)
    torch.cuda.synchronize()
    assert torch.equal(hi3, hi2) and torch.equal(lo3, lo2)
    assert torch.equal(aggr3, aggr) and torch.equal(aggr4, aggr)      # deterministic aggregation: bitwise


def test_processor_image_remainder_path_matches_row_path(golden_dir):
    """The whole 6-step loop, forward and backward, with MGN_EDGE_FWD_HL on and off."""
    from graph_physics_b200 import graph as G, ops, processor
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n1 = z["points"].shape[0]
    ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, 3)
    N, E = n1 * 3, ei.shape[1]
    K = 6
    p = mo.init_params(K, hidden=128, seed=3)
    blocks = [[p[f"processor_list.{k}.{n}"].cuda().requires_grad_(True) for n in processor.PARAM_NAMES]
              for k in range(K)]
    x0, e0, gx = processor_inputs(N, E, 128, seed=1)
    res = []
    for hl in (False, True):
        old = ops.EDGE_FWD_HL
        ops.EDGE_FWD_HL = hl
        try:
            xd = x0.cuda().requires_grad_(True)
            ed = e0.cuda().requires_grad_(True)
            xk, ek = processor.run_processor(xd, ei, ed, blocks)
            loss = (xk * gx.cuda()).sum() + (ek * ek).sum() * 1e-4
            grads = torch.autograd.grad(loss, [xd, ed] + [t for blk in blocks for t in blk])
            torch.cuda.synchronize()
            res.append((xk.detach(), ek.detach(), grads))
        finally:
            ops.EDGE_FWD_HL = old
    (x_a, e_a, g_a), (x_b, e_b, g_b) = res
    assert mo.rel_l2(x_b, x_a) < 1e-3 and mo.rel_l2(e_b, e_a) < 1e-3    # six steps of last-ulp differences
    for ga, gb in zip(g_a, g_b):
        assert mo.rel_l2(gb, ga) < 3e-2      # ReLU-mask flips of the last-ulp differences


@pytest.mark.parametrize("hub_degree", [5, 40, 300, 1000])
def test_deterministic_aggregation_hubs_and_isolated_nodes(hub_degree):
    """Receiver runs longer than a warp's 32 rows, a tile's 128 rows and several tiles (a hub node), nodes without
    in-edges (their aggr row stays zero), a ragged last tile: the ordered segmented sum (plain stores + slot carries,
    csrc/edge_fwd_hl.cu aggr_fixup_kernel) of both edge kernels against index_add_ in fp64, and bitwise against
    itself."""
    from graph_physics_b200 import graph as G, ops
    dev = "cuda"
    g = torch.Generator().manual_seed(hub_degree)
    N = 700
    rnd = torch.randint(0, N // 2, (2, 3000), generator=g)                      # nodes >= N/2 get no random in-edges
    hubs = torch.stack([torch.randint(0, N, (hub_degree,), generator=g), torch.full((hub_degree,), 123)])
    hubs2 = torch.stack([torch.randint(0, N, (hub_degree // 2 + 1,), generator=g), torch.full((hub_degree // 2 + 1,), 124)])
    ei = torch.cat([rnd, hubs, hubs2], dim=1).to(dev)
    E = ei.shape[1]
    csr = G.build_csr(ei, N)
    torch.manual_seed(7)
    ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    b = [torch.randn(128, device=dev) * 0.05 for _ in range(4)]
    sc = 1.0 + 0.1 * torch.randn(128, device=dev)
    packed = ops.pack_block(ew, nw)
    x = torch.randn(N, 128, device=dev)
    e = torch.randn(E, 128, device=dev)            # CSR order
    P = ops.p_fwd(x, packed, b[0])
    hi, lo = ops.e16_image(E, dev), ops.e16_image(E, dev)
    e1, aggr1 = ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e16_next=hi,
                             e16lo_next=lo)
    ref1 = torch.zeros(N, 128, dtype=torch.float64, device=dev).index_add_(0, csr.receivers.long(), (e1 - e).double())
    assert float((aggr1.double() - ref1).abs().max()) <= 2e-5 * float(ref1.abs().max())
    assert float(aggr1[N // 2:].abs().max()) == 0.0           # no in-edges: the row keeps its zero
    e2 = torch.empty_like(e)
    aggr2 = ops.edge_fwd_hl(hi, lo, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e_out=e2)
    ref2 = torch.zeros(N, 128, dtype=torch.float64, device=dev).index_add_(0, csr.receivers.long(), (e2 - e1).double())
    assert float((aggr2.double() - ref2).abs().max()) <= 2e-5 * float(ref2.abs().max())
    isolated = torch.ones(N, dtype=torch.bool, device=dev)
    isolated[csr.receivers.long()] = False
    assert int(isolated.sum()) > 0 and float(aggr2[isolated].abs().max()) == 0.0
    # bitwise repeatable
    # ... also when the CSR row pointers are given: no memset then, the fix-up kernel zeroes the rows of isolated nodes
    # (the output buffers are poisoned first)
    poison = lambda: torch.full((N, 128), float("nan"), device=dev)
    aggr2b = ops.edge_fwd_hl(hi, lo, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, aggr=poison(),
                             rowptr=csr.rowptr)
    _, aggr1b = ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, aggr=poison(),
                             rowptr=csr.rowptr)
    torch.cuda.synchronize()
    assert torch.equal(aggr2, aggr2b) and torch.equal(aggr1, aggr1b)


def test_pack_blocks_equals_pack_block():
    """mgn_pack_blocks (all blocks, one launch) writes the same bytes as K calls of mgn_pack_block."""
    from graph_physics_b200 import ops
    dev = "cuda"
    torch.manual_seed(3)
    K = 5
    blocks = []
    for _ in range(K):
        ew = [torch.randn(128, 384, device=dev)] + [torch.randn(128, 128, device=dev) for _ in range(3)]
        nw = [torch.randn(128, 256, device=dev)] + [torch.randn(128, 128, device=dev) for _ in range(3)]
        blocks.append(ew + nw)
    many = ops.pack_blocks(blocks)
    for k in range(K):
        one = ops.pack_block(blocks[k][:4], blocks[k][4:])
        assert torch.equal(many[k], one), k


def test_tma_gather4_path_matches_cp_async_path():
    """MGN_PJ_TMA=1 gathers the P_j[sender] rows with cp.async.bulk.tensor.2d.tile::gather4 (TMA) instead of cp.async.
    The switch is read once per process, so the image + remainder edge-kernel tests of this file and the saved-activation
    backward test are re-run in a child process with it set (they compare against kernels that do not use it)."""
    import subprocess
    import sys
    if os.environ.get("MGN_PJ_TMA_CHILD"):
        pytest.skip("already inside the child run")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, MGN_PJ_TMA="1", MGN_PJ_TMA_CHILD="1")
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "tests/test_gpu_fwd_step.py", "tests/test_gpu_bwd_step.py",
                        "-k", "image or remainder or hl or deterministic or saved"], cwd=root, env=env, capture_output=True,
                       text=True, timeout=280)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout
