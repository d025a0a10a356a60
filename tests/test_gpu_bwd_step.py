"""GPU: backward of one message-passing step (fused sm_100a kernels called through the C ABI)
vs autograd of the CPU oracle.

BASELINE.json states no gradient tolerance.  Two are used (DESIGN.md "Numerics"):
 * TOL_GRAD = 5e-3 relative L2 against the oracle evaluated in fp64 WITH the kernels' operand
   rounding points (oracle `emulate_fp16=True`) and differentiated through the SAME ReLU active
   set as the device (masks = [h > 0] of the hidden activations the forward kernels computed):
   this isolates the backward kernels' own error.
 * TOL_GRAD_EXACT = 0.10 against the exact fp64 oracle.  A ReLU network's gradient is
   discontinuous in the pre-activations: any 10-bit-mantissa forward (fp16 here, TF32 in the
   reference, train.py:28) flips the ReLU mask of the ~4e-4 fraction of units whose
   pre-activation is within rounding error of 0, which moves the gradient by sqrt(fraction)
   ~ 1.5-3.5 % per MLP although the forward agrees to 3e-4.  The fp64 emulation alone (no GPU)
   shows the same deviation (tests/test_oracle_golden.py::test_relu_mask_sensitivity)."""
import os

import numpy as np
import pytest
import torch

from oracle import mgn_oracle as mo
from oracle.synth import processor_inputs

pytestmark = pytest.mark.gpu
TOL_GRAD = 5e-3
TOL_GRAD_EXACT = 0.10


def _step_fwd_bwd(ei_np, n, x0, e0, gx1, ge1, p, use_rowmap=True):
    from graph_physics_b200 import graph as G, ops
    dev = "cuda"
    ei = torch.as_tensor(ei_np, dtype=torch.int64, device=dev)
    csr = G.build_csr(ei, n)
    pre = "processor_list.0."
    g = {k: v.to(dev) for k, v in p.items()}
    ew = [g[pre + f"edge_block.{i}.weight"] for i in (0, 2, 4, 6)]
    nw = [g[pre + f"node_block.{i}.weight"] for i in (0, 2, 4, 6)]
    eb = [g[pre + f"edge_block.{i}.bias"] for i in (0, 2, 4, 6)]
    nb = [g[pre + f"node_block.{i}.bias"] for i in (0, 2, 4, 6)]
    esc, nsc = g[pre + "edge_block.7.scale"], g[pre + "node_block.7.scale"]
    packed = ops.pack_block(ew, nw)
    xd, ed = x0.to(dev), e0.to(dev)
    perm = csr.perm.long()
    if use_rowmap:
        e_in, rowmap = ed, csr.perm
        ge = None if ge1 is None else ge1.to(dev)
    else:
        e_in, rowmap = ed.index_select(0, perm).contiguous(), None
        ge = None if ge1 is None else ge1.to(dev).index_select(0, perm).contiguous()
    P = ops.p_fwd(xd, packed, eb[0])
    E = ed.shape[0]
    esaves = tuple(torch.zeros((E, 128), dtype=torch.float16, device=dev) for _ in range(3))
    # the processor's path (CSR-ordered edges) hands the backward the forward kernel's fp16 tile images of e;
    # with a rowmap the backward converts the fp32 rows itself
    e16 = None if use_rowmap else ops.e16_image(E, dev)
    e_out, aggr = ops.edge_fwd(e_in, rowmap, P, csr.senders, csr.receivers, n, packed, eb[1], eb[2], eb[3], esc,
                               saves=esaves, e16_out=e16)
    saves = ops.node_saves(n, dev)
    x_out = ops.node_fwd(xd, aggr, packed, nb[0], nb[1], nb[2], nb[3], nsc, saves=saves)
    grad = torch.zeros(ops.GRAD_BLOCK_FLOATS, dtype=torch.float32, device=dev)
    g_x, g_aggr, dz1 = ops.node_bwd(gx1.to(dev), saves, packed, nb[3], nsc, grad)
    ops.node_wgrad1(dz1, xd, aggr, grad)
    ge_before = None if ge is None else ge.clone()
    g_e, dz16 = ops.edge_bwd(None if e16 is not None else e_in, rowmap, ge, g_aggr, P, csr.senders, csr.receivers, n,
                           packed, eb[1], eb[2], eb[3], esc, grad, e16=e16)
    if ge is not None:
        assert torch.equal(ge, ge_before)       # not in place unless asked
    if e16 is None:      # the layer-1 weight gradient needs the images of e: build them with a second forward call
        e16 = ops.e16_image(E, dev)
        ops.edge_fwd(e_in, rowmap, P, csr.senders, csr.receivers, n, packed, eb[1], eb[2], eb[3], esc, e16_out=e16)
    ops.edge_wgrad1(dz16, e16, E, grad)
    ops.p_bwd(ops.dp_gather(dz16, csr), xd, g_x, packed, grad)
    torch.cuda.synchronize()
    if not use_rowmap:
        back = torch.empty_like(g_e)
        back[perm] = g_e
        g_e = back
    emasks = []
    for hs in esaves:      # CSR position -> caller edge order
        m = torch.empty_like(hs, dtype=torch.float64)
        m[perm] = (hs > 0).double()
        emasks.append(m.cpu())
    nmasks = [(ops.image_rows(hs, n) > 0).double().cpu() for hs in saves]
    return (g_x.cpu(), g_e.cpu(), {k: v.cpu() for k, v in ops.split_grad_block(grad).items()}, emasks, nmasks)


def _oracle(ei_np, x0, e0, gx1, ge1, p, emulate, emasks=None, nmasks=None):
    p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}
    x = x0.double().requires_grad_(True)
    e = e0.double().requires_grad_(True)
    x1, e1 = mo.graph_net_block(x, torch.as_tensor(ei_np, dtype=torch.int64), e, p64, "processor_list.0",
                                emulate_fp16=emulate, edge_masks=emasks if emulate else None,
                                node_masks=nmasks if emulate else None)
    loss = (x1 * gx1.double()).sum()
    if ge1 is not None:
        loss = loss + (e1 * ge1.double()).sum()
    loss.backward()
    return x.grad, e.grad, {k[len("processor_list.0."):]: v.grad for k, v in p64.items()}


def _check(ei, n, e, seed, with_ge, use_rowmap=True):
    p = mo.init_params(1, hidden=128, seed=seed)
    x0, e0, _ = processor_inputs(n, e, 128, seed=seed + 1)
    g = torch.Generator().manual_seed(seed + 7)
    gx1 = torch.randn(n, 128, generator=g)
    ge1 = torch.randn(e, 128, generator=g) * 0.3 if with_ge else None
    g_x, g_e, grads, emasks, nmasks = _step_fwd_bwd(ei, n, x0, e0, gx1, ge1, p, use_rowmap)
    out = {}
    for emulate, tol in ((True, TOL_GRAD), (False, TOL_GRAD_EXACT)):
        rx, re, rgrads = _oracle(ei, x0, e0, gx1, ge1, p, emulate, emasks, nmasks)
        errs = {"g_x": mo.rel_l2(g_x, rx), "g_e": mo.rel_l2(g_e, re)}
        for k, v in grads.items():
            errs[k] = mo.rel_l2(v, rgrads[k])
        bad = {k: v for k, v in errs.items() if not v < tol}
        assert not bad, f"emulate={emulate}: rel-L2 over {tol}: {bad}\nall: {errs}"
        out[emulate] = errs
    print("max rel-L2 vs fp16-operand oracle %.2e, vs exact oracle %.2e" %
          (max(out[True].values()), max(out[False].values())))
    return out


@pytest.mark.parametrize("with_ge,use_rowmap", [(True, True), (False, True), (True, False)])
def test_cylinder_step_bwd(golden_dir, with_ge, use_rowmap):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    ei = z["edge_index"].astype(np.int64)
    _check(ei, 1923, ei.shape[1], 0, with_ge, use_rowmap)


@pytest.mark.parametrize("n,e", [(4, 8), (5, 1), (130, 129), (300, 1000), (1000, 257), (700, 40000)])
def test_ragged_random_graphs_bwd(n, e):
    g = torch.Generator().manual_seed(n * 1000 + e)
    ei = torch.randint(0, n, (2, e), generator=g).numpy()
    _check(ei, n, e, 3, True)


def test_processor_loop_edge_backward_equals_general_path(golden_dir):
    """edge_bwd_kernel<true> (no rowmap, upstream gradient and e images given: g built from a bulk-loaded upstream tile,
    three weight slots) against edge_bwd_kernel<false> forced by an identity rowmap, same inputs: same products on the
    same operands -> identical g_e and dz1 images; weight gradients equal up to the order of the cross-CTA reductions."""
    import numpy as np
    from graph_physics_b200 import graph as G, ops
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n1 = z["points"].shape[0]
    ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, 5)
    N, E = n1 * 5, ei.shape[1]
    assert E % 128 != 0
    csr = G.build_csr(ei, N)
    dev = "cuda"
    torch.manual_seed(11)
    ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    b = [torch.randn(128, device=dev) * 0.05 for _ in range(4)]
    sc = 1.0 + 0.1 * torch.randn(128, device=dev)
    packed = ops.pack_block(ew, nw)
    x = torch.randn(N, 128, device=dev)
    e = torch.randn(E, 128, device=dev)
    P = ops.p_fwd(x, packed, b[0])
    e16 = ops.e16_image(E, dev)
    ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e16_out=e16)
    ge = torch.randn(E, 128, device=dev)
    g_aggr = torch.randn(N, 128, device=dev)
    ident = torch.arange(E, dtype=torch.int32, device=dev)
    outs = []
    for rowmap in (None, ident):
        grad = torch.zeros(ops.GRAD_BLOCK_FLOATS, device=dev)
        dz16 = ops.e16_image(E, dev).zero_()
        g_e, _ = ops.edge_bwd(None, rowmap, ge, g_aggr, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc,
                              grad, dz16=dz16, e16=e16)
        torch.cuda.synchronize()
        outs.append((g_e, ops.image_rows(dz16, E), grad))
    (ge_f, dz_f, gr_f), (ge_g, dz_g, gr_g) = outs
    assert torch.equal(ge_f, ge_g)
    assert torch.equal(dz_f, dz_g)
    assert float((gr_f - gr_g).abs().max()) <= 1e-5 * float(gr_g.abs().max())
    # in place (what the processor loop does): same result
    ge_ip = ge.clone()
    grad = torch.zeros(ops.GRAD_BLOCK_FLOATS, device=dev)
    ops.edge_bwd(None, None, ge_ip, g_aggr, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, grad,
                 dz16=ops.e16_image(E, dev), e16=e16, inplace=True)
    torch.cuda.synchronize()
    assert torch.equal(ge_ip, ge_f)


@pytest.mark.parametrize("copies", [1, 5])
def test_edge_backward_from_saved_activations_equals_recompute(golden_dir, copies):
    """mgn_edge_bwd_saved (starts from the hidden activations mgn_edge_fwd_hl kept, g_e in place by bulk reduce-add)
    against mgn_edge_bwd's processor-loop path (recomputes them from the e image and P) on the same step: the saved
    activations are the bits the recompute produces, so dz1 images and g_e must agree to fp32 rounding of one add, the
    weight gradients to the order of the cross-CTA reductions."""
    import numpy as np
    from graph_physics_b200 import graph as G, ops
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n1 = z["points"].shape[0]
    ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, copies)
    N, E = n1 * copies, ei.shape[1]
    csr = G.build_csr(ei, N)
    dev = "cuda"
    torch.manual_seed(13 + copies)
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
    # the step under test: e_in = hi + lo; forward with and without the saves gives the same outputs
    hs = tuple(ops.e16_image(E, dev).zero_() for _ in range(3))
    hi2, lo2 = ops.e16_image(E, dev), ops.e16_image(E, dev)
    aggr_s = ops.edge_fwd_hl(hi, lo, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, hi_out=hi2,
                             lo_out=lo2, h_saves=hs, rowptr=csr.rowptr)
    hi3, lo3 = ops.e16_image(E, dev), ops.e16_image(E, dev)
    aggr_n = ops.edge_fwd_hl(hi, lo, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, hi_out=hi3,
                             lo_out=lo3, rowptr=csr.rowptr)
    torch.cuda.synchronize()
    assert torch.equal(hi2, hi3) and torch.equal(lo2, lo3) and torch.equal(aggr_s, aggr_n)
    assert all(float(ops.image_rows(h, E).abs().max()) > 0 for h in hs)
    ge = torch.randn(E, 128, device=dev)
    g_aggr = torch.randn(N, 128, device=dev)
    # reference: recompute path (FAST: no rowmap, upstream given, e image = hi)
    grad_r = torch.zeros(ops.GRAD_BLOCK_FLOATS, device=dev)
    dz_r = ops.e16_image(E, dev).zero_()
    ge_r, _ = ops.edge_bwd(None, None, ge, g_aggr, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc,
                           grad_r, dz16=dz_r, e16=hi)
    # saved path, in place, twice in a row on fresh copies (second call: stale shared-memory-independent result)
    for _ in range(2):
        grad_s = torch.zeros(ops.GRAD_BLOCK_FLOATS, device=dev)
        dz_s = ops.e16_image(E, dev).zero_()
        ge_s = ge.clone()
        ops.edge_bwd_saved(hs, ge_s, g_aggr, csr.receivers, N, packed, b[3], sc, grad_s, dz16=dz_s)
        torch.cuda.synchronize()
        assert torch.equal(ops.image_rows(dz_s, E), ops.image_rows(dz_r, E))
        assert float((ge_s - ge_r).abs().max()) <= 1e-6 * float(ge_r.abs().max())
        assert float((grad_s - grad_r).abs().max()) <= 1e-5 * float(grad_r.abs().max())


@pytest.mark.parametrize("copies", [1, 32])
def test_chained_node_wgrad1_equals_stream_order(golden_dir, copies):
    """Tile-level chaining (csrc/mlp_common.cuh): node_bwd publishes a ready flag per tile, node_wgrad1 -- launched
    directly after it -- takes tiles from a counter and waits for the flags.  Same dz1 rows, so dW1' must agree with
    the stream-ordered pair up to the order of the cross-CTA reduce-adds (fp32, different tile-to-CTA assignment:
    tolerance 1e-5 of the largest entry); the chain state is back at zero afterwards; and a second pair on the same
    state works (the flags and the counter reset themselves)."""
    from graph_physics_b200 import graph as G, ops
    dev = "cuda"
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n1 = z["points"].shape[0]
    N = n1 * copies
    torch.manual_seed(5 + copies)
    ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nb = [torch.randn(128, device=dev) * 0.05 for _ in range(4)]
    nsc = 1.0 + 0.1 * torch.randn(128, device=dev)
    packed = ops.pack_block(ew, nw)
    x, aggr, gx = (torch.randn(N, 128, device=dev) for _ in range(3))
    saves = ops.node_saves(N, dev)
    ops.node_fwd(x, aggr, packed, nb[0], nb[1], nb[2], nb[3], nsc, saves=saves)

    def pair(chain):
        grad = torch.zeros(ops.GRAD_BLOCK_FLOATS, dtype=torch.float32, device=dev)
        g_x, g_aggr, dz1 = ops.node_bwd(gx, saves, packed, nb[3], nsc, grad, chain=chain)
        ops.node_wgrad1(dz1, x, aggr, grad, chain=chain)
        torch.cuda.synchronize()
        return grad, g_x, g_aggr, dz1

    ref = pair(None)
    chain = ops.chain_state(N, dev)
    for _ in range(3):
        got = pair(chain)
        assert int(chain.abs().sum()) == 0
        for a, b in zip(got[1:], ref[1:]):
            assert torch.equal(a, b)
        w_ref, w_got = ops.split_grad_block(ref[0])["node_block.0.weight"], ops.split_grad_block(got[0])["node_block.0.weight"]
        assert float((w_got - w_ref).abs().max()) <= 1e-5 * float(w_ref.abs().max())
        rest_ref, rest_got = ref[0].clone(), got[0].clone()
        ops.split_grad_block(rest_ref)["node_block.0.weight"].zero_()
        ops.split_grad_block(rest_got)["node_block.0.weight"].zero_()
        assert float((rest_got - rest_ref).abs().max()) <= 1e-5 * float(rest_ref.abs().max())


def test_chained_entry_points_reject_half_a_chain():
    """ready_in and sched of a chained consumer go together: one without the other is an argument error (code < 0 ->
    ValueError), not a launch that would wait for flags nobody counts."""
    from graph_physics_b200 import _lib as L, ops
    L.require_device()
    dev = "cuda"
    N = 300
    x, aggr = torch.randn(N, 128, device=dev), torch.randn(N, 128, device=dev)
    dz1 = torch.zeros((N, 128), dtype=torch.float16, device=dev)
    grad = torch.zeros(ops.GRAD_BLOCK_FLOATS, dtype=torch.float32, device=dev)
    flags = ops.chain_state(N, dev)
    with pytest.raises(ValueError, match="go together"):
        L.check(L.lib().mgn_node_wgrad1c(L.i64(N), L.p(dz1), L.p(x), L.p(aggr), L.p(grad), L.p(flags), L.p(None),
                                         L.stream()))
    ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    packed = ops.pack_block(ew, nw)
    P = torch.empty((N, 256), dtype=torch.float16, device=dev)
    b1 = torch.zeros(128, device=dev)
    with pytest.raises(ValueError, match="go together"):
        L.check(L.lib().mgn_p_fwd_c(L.i64(N), L.p(x), L.p(packed), L.p(b1), L.p(P), L.p(None), L.p(flags[3:]),
                                    L.stream()))
    torch.cuda.synchronize()
    assert int(flags.abs().sum()) == 0
