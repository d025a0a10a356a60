"""The MeshGraphNet processor loop (`for block in self.processor_list`,
graphphysics/models/processors.py:130-131) as ONE autograd node over the fused sm_100a
kernels of libmgn_b200.so.

Per step k (GraphNetBlock.forward, graphphysics/models/layers.py:667-701):
  forward   P_k = p_fwd(x_k) ; e_{k+1}, aggr_k = edge_fwd(e_k, P_k) ; x_{k+1} = node_fwd(x_k, aggr_k)
  backward  node_bwd -> g_aggr, g_x (direct) ; node_wgrad1 ; edge_bwd -> g_e, dz1 images ; edge_wgrad1 ;
            dp_gather -> dP ; p_bwd -> g_x += ...
Saved per step: x_k (the residual stream itself), e_k as the fp16 tile images the edge kernel
built for its first product (the fp32 edge stream is NOT kept), aggr_k, P_k (fp16) and the node
MLP's three hidden activations (fp16).  The edge MLP's hidden activations are recomputed.
The edge-gradient stream g_e is one fp32 buffer updated in place by every edge_bwd.

Edge latents run in receiver-sorted (CSR) order inside the loop; they are permuted once on
entry and once on exit.  Gradients are multiplied by a power of two on entry to the backward
sweep (and divided on exit) so that the fp16 tensor-core operands keep their precision; the
sweep is linear in the gradient, so this is exact up to fp32 rounding.

There is no CPU / eager fallback: everything here raises unless the tensors are on an
sm_100 device and the library is built.
"""
from __future__ import annotations

import collections
import weakref
from typing import List, Optional, Sequence

import torch

from . import _lib as L
from . import graph as G
from . import ops

H = ops.H
PARAM_NAMES = tuple(n for n, _ in ops.GRAD_LAYOUT)          # 18 per block, state-dict order
_E_W = (0, 2, 4, 6)
_E_B = (1, 3, 5, 7)
_N_W = (9, 11, 13, 15)
_N_B = (10, 12, 14, 16)

# ---- instrumentation used by bench.py (kernel launch counter, per-kernel CUDA events)
LAUNCHES = 0
EVENTS: Optional[dict] = None      # name -> list of (start, end) torch.cuda.Event pairs when enabled
# data-parallel hook (graph_physics_b200.parallel): called as GRAD_SYNC(flat_block_grad, k) right after
# block k's backward has been enqueued, so the all-reduce overlaps the remaining blocks' backward
GRAD_SYNC = None
# the [K, GRAD_BLOCK_FLOATS] buffer the last backward sweep wrote: the parameters' .grad tensors are views of it
# (autograd keeps the views it is handed), so one collective on this buffer reduces every block's gradients
LAST_GRAD_BLOCKS: Optional[torch.Tensor] = None


def _timed(name, fn, *a, **kw):
    global LAUNCHES
    LAUNCHES += 1
    if EVENTS is not None and name in EVENTS:
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        r = fn(*a, **kw)
        e.record()
        EVENTS[name].append((s, e))
        return r
    return fn(*a, **kw)


_csr_cache: "collections.OrderedDict" = collections.OrderedDict()      # least recently used first
CSR_CACHE_ENTRIES = 64


def _csr_cache_put(edge_index: torch.Tensor, num_nodes: int, csr: G.CSRGraph) -> None:
    """Entries whose tensor is gone are dropped first, then the least recently used ones: a stream of short-lived batch
    tensors (batch.collate registers one per batch) must not push out the CSRs of the trajectories they are built
    from."""
    for k in [k for k, v in _csr_cache.items() if v[0]() is None]:
        del _csr_cache[k]
    _csr_cache[id(edge_index)] = (weakref.ref(edge_index), edge_index._version, int(num_nodes), csr)
    _csr_cache.move_to_end(id(edge_index))
    while len(_csr_cache) > CSR_CACHE_ENTRIES:
        _csr_cache.popitem(last=False)


def csr_for(edge_index: torch.Tensor, num_nodes: int) -> G.CSRGraph:
    """Receiver-sorted CSR of `edge_index`, cached per tensor OBJECT (weak reference + version counter): the topology
    is constant along a trajectory and callers keep passing the same tensor (the reference caches K-hop edges the
    same way, graphphysics/dataset/dataset.py:149-168).  A different tensor object is never a hit, even if the
    allocator handed it the address of a freed one."""
    key = id(edge_index)
    hit = _csr_cache.get(key)
    if hit is not None:
        ref, version, n, csr = hit
        if ref() is edge_index and version == edge_index._version and n == int(num_nodes):
            _csr_cache.move_to_end(key)
            return csr
    csr = G.build_csr(edge_index, num_nodes)
    _csr_cache_put(edge_index, num_nodes, csr)
    return csr


def register_csr(edge_index: torch.Tensor, num_nodes: int, csr: G.CSRGraph) -> None:
    """Hands `csr_for` a CSR that was assembled elsewhere for this tensor (batch.collate: the concatenation of the
    samples' cached CSRs), so the sort is not launched for it."""
    if csr.num_nodes != int(num_nodes) or csr.num_edges != int(edge_index.shape[1]):
        raise ValueError("register_csr: the CSR does not match edge_index / num_nodes")
    _csr_cache_put(edge_index, num_nodes, csr)


def _check_inputs(x, e, params, num_steps):
    L.require_device(x)
    if x.dtype != torch.float32 or e.dtype != torch.float32:
        raise ValueError("graph_physics_b200 processor takes fp32 latents")
    if x.dim() != 2 or x.shape[1] != H or e.dim() != 2 or e.shape[1] != H:
        raise ValueError(f"graph_physics_b200 processor supports hidden_size={H} only; got x {tuple(x.shape)}, "
                         f"edge_attr {tuple(e.shape)}")
    if len(params) != 18 * num_steps:
        raise ValueError("expected 18 parameters per GraphNetBlock (nb_of_layers=4, layer_norm=True)")


class _ProcessorFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x0, e0, csr, num_steps, need_edges, norm_dim, *params):
        ctx.set_materialize_grads(False)
        K = int(num_steps)
        nd = int(norm_dim)
        _check_inputs(x0, e0, params, K)
        dev = x0.device
        N, E = x0.shape[0], e0.shape[0]
        need_grad = any(ctx.needs_input_grad)   # False under torch.no_grad() / all-frozen inputs
        prm = [p.detach().contiguous() for p in params]
        packed = torch.empty((K, ops.packed_block_bytes()), dtype=torch.uint8, device=dev)
        _timed("pack", ops.pack_blocks, [[prm[18 * k + i] for i in _E_W + _N_W] for k in range(K)], out=packed)
        x = x0.detach().contiguous()
        # edges run in receiver-sorted (CSR) order inside the loop; the first edge kernel gathers its input rows
        # through the permutation (no separate pass over the edge stream)
        e = e0.detach().contiguous()
        rowmap0 = None if csr.identity else csr.perm
        xs, aggrs, Ps, saves, e16s, esaves = [x], [], [], [], [], []
        hl = ops.EDGE_FWD_HL and E > 0
        # the fp32 edge stream ping-pongs; in image + remainder form (hl) only step 0 reads fp32 rows (the caller's)
        # and only the last step writes them, and only if the caller wants the edge latents back
        e_bufs = [torch.empty_like(e) for _ in range(min(K, 2))] if not hl else None
        e_last = torch.empty_like(e) if (hl and (need_edges or K == 1)) else None
        lo_bufs = [ops.e16_image(E, dev) for _ in range(min(K - 1, 2))] if hl else None
        if not need_grad:   # inference: ping-pong buffers, nothing kept
            P = torch.empty((N, 2 * H), dtype=torch.float16, device=dev)
            aggr = torch.empty((N, H), dtype=torch.float32, device=dev)
            x_bufs = [torch.empty_like(x) for _ in range(min(K, 2))]
        # fp16 operand-tile images of the edge stream: step k reads image k (step 0 builds it while loading e_0) and
        # writes image k+1 next to e_{k+1}; training keeps them all (the edge backward recomputes from them)
        img_bufs = [ops.e16_image(E, dev) for _ in range(min(K, 2))] if not need_grad else None
        img = None
        # workspace of the deterministic aggregation (ops.carry_buffer): 8-row slots for step 0, 32-row slots after
        carry = ops.carry_buffer(E, 8, dev) if (ops.DETERMINISTIC and E > 0) else None
        chain_np = ops.chain_state(N, dev) if (ops.CHAIN and K > 1) else None    # node_fwd(k) -> p_fwd(k+1), tile by tile
        for k in range(K):
            b = prm[18 * k:18 * (k + 1)]
            if need_grad:
                P = torch.empty((N, 2 * H), dtype=torch.float16, device=dev)
                aggr = torch.empty((N, H), dtype=torch.float32, device=dev)
                x_out = torch.empty_like(x)
                sv = ops.node_saves(N, dev)
            else:
                x_out, sv = x_bufs[k & 1], None
            last = k + 1 == K
            if hl:
                e_out = e_last if last else None
            else:
                e_out = e_bufs[k & 1]
            if k == 0:
                img = ops.e16_image(E, dev) if need_grad else img_bufs[1 % len(img_bufs)]
                img_self, img_in = img, None
            else:
                img_self, img_in = None, img
            img_next = None
            if not last:
                img_next = ops.e16_image(E, dev) if need_grad else img_bufs[k & 1]
            lo_next = lo_bufs[k & 1] if (hl and not last) else None
            _timed("p_fwd", ops.p_fwd, x, packed[k], b[1], out=P, chain=chain_np if k > 0 else None)
            esv = None
            if hl and k > 0:
                # training: the edge MLP's hidden activations are kept as tile images (the edge backward of this step
                # then starts at layer 4 instead of recomputing the block)
                if need_grad and ops.SAVE_EDGE_HIDDEN:
                    esv = tuple(ops.e16_image(E, dev) for _ in range(3))
                _timed("edge_fwd", ops.edge_fwd_hl, img_in, lo_bufs[(k - 1) & 1], P, csr.senders, csr.receivers, N,
                       packed[k], b[3], b[5], b[7], b[8], hi_out=img_next, lo_out=lo_next, e_out=e_out, aggr=aggr,
                       carry=carry, norm_dim=nd, rowptr=csr.rowptr, h_saves=esv)
            else:
                _timed("edge_fwd", ops.edge_fwd, e, rowmap0 if k == 0 else None, P, csr.senders, csr.receivers, N,
                       packed[k], b[3], b[5], b[7], b[8], e_out=e_out, aggr=aggr, e16_out=img_self, e16_in=img_in,
                       e16_next=img_next, out_in_csr_order=True, e16lo_next=lo_next,
                       write_rows=not (hl and e_out is None), carry=carry, norm_dim=nd, rowptr=csr.rowptr)
            _timed("node_fwd", ops.node_fwd, x, aggr, packed[k], b[10], b[12], b[14], b[16], b[17], x_out=x_out,
                   saves=sv, norm_dim=nd, chain=None if last else chain_np)
            if need_grad:
                aggrs.append(aggr); Ps.append(P); saves.append(sv); xs.append(x_out); e16s.append(img)
                esaves.append(esv)
            x, img = x_out, img_next
            if e_out is not None:
                e = e_out
        if not need_edges:          # the caller discards the edge latents (EncodeProcessDecode): skip the exit permutation
            e_ret = e.new_empty((0, H))
            ctx.mark_non_differentiable(e_ret)
        elif csr.identity:
            e_ret = e
        else:
            e_ret = torch.empty_like(e)
            e_ret.index_copy_(0, csr.perm_long, e)
        if need_grad:
            ctx.csr, ctx.K, ctx.packed, ctx.prm = csr, K, packed, prm
            # the edge backward needs e_k only as the fp16 tile images: the fp32 versions are not kept
            ctx.xs, ctx.e16s, ctx.aggrs, ctx.Ps, ctx.saves, ctx.esaves = xs, e16s, aggrs, Ps, saves, esaves
            ctx.E, ctx.nd = E, nd
        return x, e_ret

    @staticmethod
    def backward(ctx, g_x, g_e):
        csr, K, packed, prm = ctx.csr, ctx.K, ctx.packed, ctx.prm
        xs, e16s = ctx.xs, ctx.e16s
        dev = xs[0].device
        N, E, nd = xs[0].shape[0], ctx.E, ctx.nd
        if g_x is None and g_e is None:
            return (None,) * (6 + 18 * K)
        # ---- power-of-two gradient scale (device-side, no host sync)
        amax = torch.zeros((), dtype=torch.float32, device=dev)
        if g_x is not None:
            amax = torch.maximum(amax, g_x.detach().abs().max())
        if g_e is not None:
            amax = torch.maximum(amax, g_e.detach().abs().max())
        s = torch.exp2(torch.floor(torch.log2(64.0 / amax.clamp_min(1e-30)))).clamp(2.0 ** -100, 2.0 ** 100)
        inv_s = 1.0 / s
        gx = (g_x.detach() * s) if g_x is not None else torch.zeros((N, H), dtype=torch.float32, device=dev)
        gx = gx.contiguous()
        if g_e is not None:
            ge = g_e.detach() if csr.identity else g_e.detach().index_select(0, csr.perm_long)
            ge = (ge * s).contiguous()
        else:
            ge = None
        global LAST_GRAD_BLOCKS
        grads = torch.zeros((K, ops.GRAD_BLOCK_FLOATS), dtype=torch.float32, device=dev)
        LAST_GRAD_BLOCKS = grads
        gx_new = torch.empty_like(gx)
        g_aggr = torch.empty((N, H), dtype=torch.float32, device=dev)
        dz1 = torch.empty((N, H), dtype=torch.float16, device=dev)
        chain_nw = ops.chain_state(N, dev) if ops.CHAIN else None     # node_bwd -> node_wgrad1, tile by tile
        chain_pn = ops.chain_state(N, dev) if (ops.CHAIN and K > 1) else None     # p_bwd(k) -> node_bwd(k-1)
        dz16 = ops.e16_image(E, dev)       # dz1 of the current step: written by edge_bwd, gathered by p_bwd
        dP = torch.empty((N, 2 * H), dtype=torch.float16, device=dev)
        if E > 0:
            csr.by_sender                  # (built once per topology)
        else:
            dP.zero_()                     # no edges: nothing flows back through the gathers
        if ge is None:      # no upstream edge gradient: the last block writes g_e, the others accumulate into it
            ge_buf, have_ge = torch.empty((E, H), dtype=torch.float32, device=dev), False
        else:
            ge_buf, have_ge = ge, True
        for k in range(K - 1, -1, -1):
            b = prm[18 * k:18 * (k + 1)]
            _timed("node_bwd", ops.node_bwd, gx, ctx.saves[k], packed[k], b[16], b[17], grads[k], g_x=gx_new,
                   g_aggr=g_aggr, dz1=dz1, norm_dim=nd, chain=chain_nw, chain_in=chain_pn if k < K - 1 else None)
            _timed("node_wgrad1", ops.node_wgrad1, dz1, xs[k], ctx.aggrs[k], grads[k], chain=chain_nw)
            if E > 0:
                if ctx.esaves[k] is not None:
                    if not have_ge:          # no upstream edge gradient: the in-place kernel adds to zeros
                        ge_buf.zero_()
                    _timed("edge_bwd", ops.edge_bwd_saved, ctx.esaves[k], ge_buf, g_aggr, csr.receivers, N, packed[k],
                           b[7], b[8], grads[k], dz16=dz16, norm_dim=nd)
                else:
                    _timed("edge_bwd", ops.edge_bwd, None, None, ge_buf if have_ge else None, g_aggr, ctx.Ps[k],
                           csr.senders, csr.receivers, N, packed[k], b[3], b[5], b[7], b[8], grads[k], g_e=ge_buf,
                           dz16=dz16, e16=e16s[k], inplace=True, norm_dim=nd)
                have_ge = True
                _timed("edge_wgrad1", ops.edge_wgrad1, dz16, e16s[k], E, grads[k])
                _timed("dp_gather", ops.dp_gather, dz16, csr, dP)
            _timed("p_bwd", ops.p_bwd, dP, xs[k], gx_new, packed[k], grads[k], chain=chain_pn if k > 0 else None)
            gx, gx_new = gx_new, gx
            if GRAD_SYNC is not None:        # the reduction of block k starts now: it needs the final values
                grads[k].mul_(inv_s)
                GRAD_SYNC(grads[k], k)
            # release what this step saved
            ctx.saves[k] = None; ctx.aggrs[k] = None; ctx.Ps[k] = None; e16s[k] = None; ctx.esaves[k] = None
            xs[k + 1] = None
        if GRAD_SYNC is None:
            grads.mul_(inv_s)                # one launch for all blocks
        if GRAD_SYNC is not None and hasattr(GRAD_SYNC, "finish"):
            # the reductions started per block must be complete (in stream order) before autograd hands the views
            # below to AccumulateGrad: with an existing .grad it would otherwise add partly reduced values
            GRAD_SYNC.finish()
        g_x0 = gx * inv_s if ctx.needs_input_grad[0] else None
        g_e0 = None
        if ctx.needs_input_grad[1]:
            ge = ge_buf * inv_s
            if csr.identity:
                g_e0 = ge
            else:
                g_e0 = torch.empty_like(ge)
                g_e0.index_copy_(0, csr.perm_long, ge)
        out: List[Optional[torch.Tensor]] = [g_x0, g_e0, None, None, None, None]
        for k in range(K):
            views = ops.split_grad_block(grads[k])
            for i, name in enumerate(PARAM_NAMES):
                out.append(views[name] if ctx.needs_input_grad[6 + 18 * k + i] else None)
        return tuple(out)


def run_processor(x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor,
                  block_params: Sequence[Sequence[torch.Tensor]], csr: Optional[G.CSRGraph] = None,
                  need_edges: bool = True):
    """x [N,d], edge_index [2,E] int64, edge_attr [E,d] (caller's edge order), d <= 128;
    block_params: per GraphNetBlock the 18 tensors in ops.GRAD_LAYOUT order AT THE KERNELS' SHAPE (hidden 128, four
    layers): for d < 128 or fewer layers `models.layers.block_param_list` zero-pads / inserts identity layers.
    Latents narrower than 128 are zero-padded here and the result is cut back; RMSNorm divides by sqrt(d).
    Returns (x_K, edge_attr_K) with edge_attr_K in the caller's edge order (an empty tensor if not need_edges)."""
    L.require_device(x)
    if csr is None:
        csr = csr_for(edge_index, x.shape[0])
    if csr.num_edges != edge_attr.shape[0] or csr.num_nodes != x.shape[0]:
        raise ValueError("edge_index / edge_attr / x sizes disagree")
    d = int(x.shape[1])
    if edge_attr.shape[1] != d or not (0 < d <= H):
        raise ValueError(f"graph_physics_b200 processor supports hidden_size <= {H} with equal node / edge widths; got "
                         f"x {tuple(x.shape)}, edge_attr {tuple(edge_attr.shape)}")
    if d < H:
        x = torch.nn.functional.pad(x, (0, H - d))
        edge_attr = torch.nn.functional.pad(edge_attr, (0, H - d))
    flat = [t for blk in block_params for t in blk]
    xo, eo = _ProcessorFn.apply(x, edge_attr, csr, len(block_params), bool(need_edges), d, *flat)
    if d < H:
        xo = xo[:, :d]
        eo = eo[:, :d] if need_edges else eo
    return xo, eo


class CapturedStep:
    """Capture `fn` (a whole training or inference step: module forward, backward, in-place gradient hooks) into a
    CUDA graph after `warmup` eager runs and replay it on every call.  Every launch of this package is stream-ordered
    and never reads device data on the host, so a step captures as is; tensors `fn` reads must be updated in place
    between replays, and the outputs / .grad tensors of the capture are the ones every replay writes."""

    def __init__(self, fn, warmup: int = 3):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        # thread_local: other threads (autograd workers, the NCCL watchdog) may keep making CUDA calls
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            self.result = fn()

    def __call__(self):
        self.graph.replay()
        return self.result
