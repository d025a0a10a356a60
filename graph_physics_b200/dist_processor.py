"""The processor loop on ONE mesh cut over several GPUs (BASELINE.json config 5).

Every rank runs the same fused kernels as `processor.py` on its partition (`partition.Partition`): its owned nodes,
the halo copies of remote senders, and every edge whose receiver it owns.  Per message-passing step:

  forward   P[owned] = p_fwd(x_k) ; halo rows of P's sender half <- owners (all-to-all-v, fp16) ;
            e_{k+1}, aggr = edge_fwd ; x_{k+1}[owned] = node_fwd
  backward  node_bwd, node_wgrad1 on the owned rows ; edge_bwd, edge_wgrad1, dp_gather on the local edges ;
            halo rows of dP's sender half -> owners, added there (all-to-all-v, fp16) ; p_bwd on the owned rows

and once per sweep: the gradient pre-scale is agreed with a MAX all-reduce (one float), the 15 flat weight-gradient
blocks are SUMMED over ranks (every rank holds a partial sum over its edges / nodes).

`PartitionEngine` is the step-wise state machine (no communication inside): `PartitionedProcessorFn` drives it with
torch.distributed collectives as one autograd node; `run_partitions_local` drives several engines in one process
(single-GPU validation of the partition logic against `processor.run_processor`).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from . import _lib as L
from . import graph as G
from . import ops
from .partition import HaloExchange, LocalHaloExchange, Partition
from .processor import PARAM_NAMES, _E_W, _N_W, _check_inputs, _timed

H = ops.H


class PartitionEngine:
    """One rank's share of the loop.  What crosses a partition boundary is what the edge kernels gather by SENDER: the
    sender half of the per-node partial products P (fp16, 256 B per node) in the forward, and the sender half of dP in
    the backward -- half the bytes of the x / g_x rows, and no rank computes P, dP or g_x for nodes it does not own:

        fwd_p(k)   P[:owned] = p_fwd(x_k)                ; caller: halo rows of P_j <- owners (halo_send_fwd / halo_recv_fwd)
        fwd_rest(k) edge update, aggregation, node update on the owned nodes
        bwd_a(k)   node_bwd .. dp_gather -> dP (owned + halo) ; caller: owners' dP_j += halo rows (halo_send_bwd / halo_recv_bwd)
        bwd_b(k)   p_bwd on the owned nodes
    """

    def __init__(self, part: Partition, block_params: Sequence[Sequence[torch.Tensor]], need_grad: bool = True,
                 csr: Optional[G.CSRGraph] = None, need_edges: bool = True):
        self.part, self.K, self.need_grad = part, len(block_params), need_grad
        self.need_edges = bool(need_edges)      # False: the fp32 rows of the final edge latents are never written
        if part.num_owned == 0:
            raise ValueError("a partition must own at least one node")
        self.csr = csr if csr is not None else G.build_csr(part.edge_index, part.num_ext)
        self.prm = [[p.detach().contiguous() for p in blk] for blk in block_params]
        self.packed = None

    # ------------------------------------------------------------------ forward
    def fwd_begin(self, x_owned: torch.Tensor, e_local: torch.Tensor) -> None:
        part, K = self.part, self.K
        _check_inputs(x_owned, e_local, [t for b in self.prm for t in b], K)
        if x_owned.shape[0] != part.num_owned or e_local.shape[0] != self.csr.num_edges:
            raise ValueError("x_owned / e_local do not match the partition")
        dev = x_owned.device
        self.dev, self.E = dev, e_local.shape[0]
        self.packed = torch.empty((K, ops.packed_block_bytes()), dtype=torch.uint8, device=dev)
        _timed("pack", ops.pack_blocks, [[b[i] for i in _E_W + _N_W] for b in self.prm], out=self.packed)
        self.xs: List[Optional[torch.Tensor]] = [x_owned.detach().contiguous()]
        self.e = e_local.detach().contiguous()
        # between steps the edge stream is an fp16 image + fp16 remainder (csrc/edge_fwd_hl.cu, as in processor.py):
        # only step 0 reads fp32 rows, only the last step writes them (and only if somebody wants them)
        self.hl = ops.EDGE_FWD_HL and self.E > 0
        self.e_bufs = [torch.empty_like(self.e) for _ in range(min(K, 2))] if not self.hl else None
        self.e_last = torch.empty_like(self.e) if (self.hl and (self.need_edges or K == 1)) else None
        self.lo_bufs = [ops.e16_image(self.E, dev) for _ in range(min(K - 1, 2))] if self.hl else None
        self.aggrs, self.Ps, self.saves, self.e16s, self.esaves = [], [], [], [], []
        self.img = None
        self.carry = ops.carry_buffer(self.E, 8, dev) if (ops.DETERMINISTIC and self.E > 0) else None
        if not self.need_grad:
            self.P = torch.empty((part.num_ext, 2 * H), dtype=torch.float16, device=dev)
            self.aggr = torch.empty((part.num_ext, H), dtype=torch.float32, device=dev)
            self.img_bufs = [ops.e16_image(self.E, dev) for _ in range(min(K, 2))]

    def x_cur(self, k: int) -> torch.Tensor:
        return self.xs[k] if self.need_grad else self.xs[-1]

    def fwd_p(self, k: int) -> None:
        """P of the owned nodes; the halo rows' sender halves come from their owners next."""
        part, dev = self.part, self.dev
        n_own, n_ext = part.num_owned, part.num_ext
        self.P_k = torch.empty((n_ext, 2 * H), dtype=torch.float16, device=dev) if self.need_grad else self.P
        _timed("p_fwd", ops.p_fwd, self.x_cur(k), self.packed[k], self.prm[k][1], out=self.P_k[:n_own])

    def halo_send_fwd(self) -> torch.Tensor:
        """[len(send_index), 128] fp16: sender halves of the owned rows the other ranks gather from."""
        p = self.part
        return self.P_k[:p.num_owned].index_select(0, p.send_index)[:, H:].contiguous()

    def halo_recv_fwd(self, rows: torch.Tensor) -> None:
        self.P_k[self.part.num_owned:, H:] = rows

    def fwd_rest(self, k: int) -> None:
        part, K, dev, E = self.part, self.K, self.dev, self.E
        n_own, n_ext = part.num_owned, part.num_ext
        b, csr, ng = self.prm[k], self.csr, self.need_grad
        x = self.x_cur(k)
        x_out = torch.empty((n_own, H), dtype=torch.float32, device=dev)
        P = self.P_k
        if ng:
            aggr = torch.empty((n_ext, H), dtype=torch.float32, device=dev)
            sv = ops.node_saves(n_own, dev)
        else:
            aggr, sv = self.aggr, None
        if E > 0:
            last = k + 1 == K
            e_out = (self.e_last if last else None) if self.hl else self.e_bufs[k & 1]
            if k == 0:
                self.img = ops.e16_image(E, dev) if ng else self.img_bufs[1 % len(self.img_bufs)]
                img_self, img_in = self.img, None
            else:
                img_self, img_in = None, self.img
            img_next = None
            if not last:
                img_next = ops.e16_image(E, dev) if ng else self.img_bufs[k & 1]
            lo_next = self.lo_bufs[k & 1] if (self.hl and not last) else None
            rowmap = csr.perm if (k == 0 and not csr.identity) else None
            esv = None
            if self.hl and k > 0:
                if ng and ops.SAVE_EDGE_HIDDEN:     # as processor.py: the edge backward starts from these
                    esv = tuple(ops.e16_image(E, dev) for _ in range(3))
                _timed("edge_fwd", ops.edge_fwd_hl, img_in, self.lo_bufs[(k - 1) & 1], P, csr.senders, csr.receivers,
                       n_ext, self.packed[k], b[3], b[5], b[7], b[8], hi_out=img_next, lo_out=lo_next, e_out=e_out,
                       aggr=aggr, carry=self.carry, rowptr=csr.rowptr, h_saves=esv)
            else:
                _timed("edge_fwd", ops.edge_fwd, self.e, rowmap, P, csr.senders, csr.receivers, n_ext, self.packed[k],
                       b[3], b[5], b[7], b[8], e_out=e_out, aggr=aggr, e16_out=img_self, e16_in=img_in,
                       e16_next=img_next, out_in_csr_order=True, e16lo_next=lo_next,
                       write_rows=not (self.hl and e_out is None), carry=self.carry, rowptr=csr.rowptr)
            if ng:
                self.e16s.append(self.img)
                self.esaves.append(esv)
            self.img = img_next
            if e_out is not None:
                self.e = e_out
        else:
            aggr.zero_()
            if ng:
                self.e16s.append(None)
                self.esaves.append(None)
        _timed("node_fwd", ops.node_fwd, x, aggr[:n_own], self.packed[k], b[10], b[12], b[14], b[16], b[17],
               x_out=x_out, saves=sv)
        if ng:
            self.aggrs.append(aggr); self.Ps.append(P); self.saves.append(sv); self.xs.append(x_out)
        else:
            self.xs = [x_out]

    def outputs(self, need_edges: bool = True):
        """(x_K of the owned nodes, e_K of the local edges in the caller's local edge order)."""
        x = self.xs[-1]
        if not need_edges:
            return x, self.e.new_empty((0, H))
        if not self.need_edges and self.hl and self.K > 1:
            raise ValueError("this engine was built with need_edges=False: the final edge rows were not written")
        if self.csr.identity or self.E == 0:
            return x, self.e
        e_ret = torch.empty_like(self.e)
        e_ret.index_copy_(0, self.csr.perm_long, self.e)
        return x, e_ret

    # ------------------------------------------------------------------ backward
    def grad_amax(self, g_x: Optional[torch.Tensor], g_e: Optional[torch.Tensor]) -> torch.Tensor:
        amax = torch.zeros((), dtype=torch.float32, device=self.dev)
        if g_x is not None:
            amax = torch.maximum(amax, g_x.detach().abs().max())
        if g_e is not None and g_e.numel():
            amax = torch.maximum(amax, g_e.detach().abs().max())
        return amax

    @staticmethod
    def scale_from_amax(amax: torch.Tensor) -> torch.Tensor:
        """The power of two the sweep is multiplied by (same rule as processor.py); `amax` must be the maximum over
        ALL ranks, because halo gradients of different ranks are added together."""
        return torch.exp2(torch.floor(torch.log2(64.0 / amax.clamp_min(1e-30)))).clamp(2.0 ** -100, 2.0 ** 100)

    def bwd_begin(self, g_x: Optional[torch.Tensor], g_e: Optional[torch.Tensor], s: torch.Tensor) -> None:
        part, K, dev, E, csr = self.part, self.K, self.dev, self.E, self.csr
        n_own, n_ext = part.num_owned, part.num_ext
        self.s, self.inv_s = s, 1.0 / s
        self.gx = (g_x.detach() * s).contiguous() if g_x is not None else \
            torch.zeros((n_own, H), dtype=torch.float32, device=dev)
        self.gx_new = torch.empty_like(self.gx)
        if g_e is not None and E > 0:
            ge = g_e.detach() if csr.identity else g_e.detach().index_select(0, csr.perm_long)
            self.ge_buf, self.have_ge = (ge * s).contiguous(), True
        else:
            self.ge_buf, self.have_ge = torch.empty((E, H), dtype=torch.float32, device=dev), False
        self.grads = torch.zeros((K, ops.GRAD_BLOCK_FLOATS), dtype=torch.float32, device=dev)
        self.g_aggr = torch.empty((n_ext, H), dtype=torch.float32, device=dev)
        self.dz1 = torch.empty((n_own, H), dtype=torch.float16, device=dev)
        self.dz16 = ops.e16_image(E, dev)
        self.dP = torch.empty((n_ext, 2 * H), dtype=torch.float16, device=dev)
        if E > 0:
            csr.by_sender
        else:
            self.dP.zero_()

    def bwd_a(self, k: int) -> None:
        """node_bwd .. dp_gather: dP of the owned AND the halo nodes (the halo rows hold what this rank's edges send)."""
        part, E, csr = self.part, self.E, self.csr
        n_own, n_ext = part.num_owned, part.num_ext
        b, pk, gk = self.prm[k], self.packed[k], self.grads[k]
        gx, gx_new = self.gx, self.gx_new
        _timed("node_bwd", ops.node_bwd, gx, self.saves[k], pk, b[16], b[17], gk, g_x=gx_new,
               g_aggr=self.g_aggr[:n_own], dz1=self.dz1)
        _timed("node_wgrad1", ops.node_wgrad1, self.dz1, self.xs[k], self.aggrs[k][:n_own], gk)
        if E > 0:
            if self.esaves[k] is not None:
                if not self.have_ge:
                    self.ge_buf.zero_()
                _timed("edge_bwd", ops.edge_bwd_saved, self.esaves[k], self.ge_buf, self.g_aggr, csr.receivers, n_ext,
                       pk, b[7], b[8], gk, dz16=self.dz16)
            else:
                _timed("edge_bwd", ops.edge_bwd, None, None, self.ge_buf if self.have_ge else None, self.g_aggr,
                       self.Ps[k], csr.senders, csr.receivers, n_ext, pk, b[3], b[5], b[7], b[8], gk, g_e=self.ge_buf,
                       dz16=self.dz16, e16=self.e16s[k], inplace=True)
            self.have_ge = True
            _timed("edge_wgrad1", ops.edge_wgrad1, self.dz16, self.e16s[k], E, gk)
            _timed("dp_gather", ops.dp_gather, self.dz16, csr, self.dP)

    def halo_send_bwd(self) -> torch.Tensor:
        """[num_halo, 128] fp16: sender halves of dP this rank holds for its halo nodes."""
        return self.dP[self.part.num_owned:, H:].contiguous()

    def halo_recv_bwd(self, rows: torch.Tensor) -> None:
        """rows [len(send_index), 128] fp16 from the other ranks: added to the owners' dP_j.  A slab cut of a grid sends
        every boundary node to ONE neighbour: `send_index` then has no duplicates and the adds have one order; a node
        that several ranks gather from receives its (two or three) fp16 contributions in arrival order."""
        p = self.part
        if rows.numel():
            self.dP[:p.num_owned, H:].index_add_(0, p.send_index, rows)

    def bwd_b(self, k: int) -> None:
        n_own = self.part.num_owned
        pk, gk = self.packed[k], self.grads[k]
        gx, gx_new = self.gx, self.gx_new
        _timed("p_bwd", ops.p_bwd, self.dP[:n_own], self.xs[k], gx_new, pk, gk)
        self.gx, self.gx_new = gx_new, gx
        gk.mul_(self.inv_s)
        self.saves[k] = None; self.aggrs[k] = None; self.Ps[k] = None; self.e16s[k] = None; self.esaves[k] = None
        self.xs[k + 1] = None

    def input_grads(self, need_gx: bool = True, need_ge: bool = True):
        """(g_x0 of the owned nodes, g_e0 of the local edges in the caller's local order); call after the halo return
        of step 0."""
        g_x0 = self.gx * self.inv_s if need_gx else None
        g_e0 = None
        if need_ge:
            ge = (self.ge_buf * self.inv_s) if self.have_ge else torch.zeros_like(self.ge_buf)
            if self.csr.identity or self.E == 0:
                g_e0 = ge
            else:
                g_e0 = torch.empty_like(ge)
                g_e0.index_copy_(0, self.csr.perm_long, ge)
        return g_x0, g_e0


# ---------------------------------------------------------------------- torch.distributed driver
class PartitionedProcessorFn(torch.autograd.Function):
    """(x_owned, e_local) -> (x_K owned, e_K local) on this rank's partition; collectives inside forward and backward."""

    @staticmethod
    def forward(ctx, x_owned, e_local, part, csr, group, num_steps, need_edges, *params):
        ctx.set_materialize_grads(False)
        K = int(num_steps)
        blocks = [params[18 * k:18 * (k + 1)] for k in range(K)]
        need_grad = any(ctx.needs_input_grad)
        eng = PartitionEngine(part, blocks, need_grad=need_grad, csr=csr, need_edges=need_edges)
        halo = HaloExchange(part, group)
        eng.fwd_begin(x_owned, e_local)
        multi = part.world > 1
        recv = torch.empty((part.num_halo, H), dtype=torch.float16, device=x_owned.device) if multi else None
        for k in range(K):
            eng.fwd_p(k)
            if multi:
                halo.rows_forward(eng.halo_send_fwd(), recv)
                eng.halo_recv_fwd(recv)
            eng.fwd_rest(k)
        x, e = eng.outputs(need_edges)
        if not need_edges:
            ctx.mark_non_differentiable(e)
        if need_grad:
            ctx.eng, ctx.halo, ctx.group, ctx.K = eng, halo, group, K
        return x, e

    @staticmethod
    def backward(ctx, g_x, g_e):
        eng, halo, K = ctx.eng, ctx.halo, ctx.K
        # every rank must take part in the collectives, also one whose outputs received no gradient
        amax = eng.grad_amax(g_x, g_e)
        multi = eng.part.world > 1
        if multi:
            dist.all_reduce(amax, op=dist.ReduceOp.MAX, group=ctx.group)
        eng.bwd_begin(g_x, g_e, eng.scale_from_amax(amax))
        recv = torch.empty((int(eng.part.send_index.numel()), H), dtype=torch.float16, device=eng.dev) if multi else None
        for k in range(K - 1, -1, -1):
            eng.bwd_a(k)
            if multi:
                halo.rows_backward(eng.halo_send_bwd(), recv)
                eng.halo_recv_bwd(recv)
            eng.bwd_b(k)
        if multi:
            dist.all_reduce(eng.grads, op=dist.ReduceOp.SUM, group=ctx.group)
        g_x0, g_e0 = eng.input_grads(ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        out: List[Optional[torch.Tensor]] = [g_x0, g_e0, None, None, None, None, None]
        for k in range(K):
            views = ops.split_grad_block(eng.grads[k])
            for i, name in enumerate(PARAM_NAMES):
                out.append(views[name] if ctx.needs_input_grad[7 + 18 * k + i] else None)
        ctx.eng = None
        return tuple(out)


def run_partitioned(x_owned: torch.Tensor, e_local: torch.Tensor, part: Partition,
                    block_params: Sequence[Sequence[torch.Tensor]], group=None, csr: Optional[G.CSRGraph] = None,
                    need_edges: bool = True):
    """This rank's share of `processor.run_processor` on a mesh cut by `partition.build_partition`.
    x_owned [num_owned,128]: latents of the owned nodes; e_local [E_loc,128]: latents of the local edges, rows in
    the order of `part.edge_ids`.  Weight gradients come back already summed over the ranks of `group`."""
    L.require_device(x_owned)
    flat = [t for blk in block_params for t in blk]
    return PartitionedProcessorFn.apply(x_owned, e_local, part, csr, group, len(block_params), bool(need_edges),
                                        *flat)


# ---------------------------------------------------------------------- single-process driver
def run_partitions_local(parts: Sequence[Partition], x: torch.Tensor, e: torch.Tensor,
                         block_params: Sequence[Sequence[torch.Tensor]], g_x: Optional[torch.Tensor] = None,
                         g_e: Optional[torch.Tensor] = None):
    """All partitions of one mesh on ONE device, stepped in lock-step with copies in place of the collectives.
    x [N,128], e [E,128], g_x / g_e: upstream gradients (global numbering).  Returns a dict with x_out, e_out and, if
    a gradient was given, g_x, g_e and grads ([K, GRAD_BLOCK_FLOATS], summed over partitions)."""
    K = len(block_params)
    need_grad = g_x is not None or g_e is not None
    engs = [PartitionEngine(p, block_params, need_grad=need_grad) for p in parts]
    halo = LocalHaloExchange(parts)
    for p, eng in zip(parts, engs):
        lo = p.node_begin
        eng.fwd_begin(x[lo:lo + p.num_owned].contiguous(), e.index_select(0, p.edge_ids))
    for k in range(K):
        for eng in engs:
            eng.fwd_p(k)
        halo.rows_forward_all([eng.P_k[:eng.part.num_owned, H:] for eng in engs],
                              [eng.P_k[eng.part.num_owned:, H:] for eng in engs])
        for eng in engs:
            eng.fwd_rest(k)
    x_out, e_out = torch.empty_like(x), torch.empty_like(e)
    for p, eng in zip(parts, engs):
        xo, eo = eng.outputs(True)
        x_out[p.node_begin:p.node_begin + p.num_owned] = xo
        e_out.index_copy_(0, p.edge_ids, eo)
    res = {"x_out": x_out, "e_out": e_out}
    if not need_grad:
        return res
    loc = []
    for p, eng in zip(parts, engs):
        lo = p.node_begin
        gxl = g_x[lo:lo + p.num_owned] if g_x is not None else None
        gel = g_e.index_select(0, p.edge_ids) if g_e is not None else None
        loc.append((gxl, gel))
    amax = torch.stack([eng.grad_amax(*l) for eng, l in zip(engs, loc)]).max()
    s = PartitionEngine.scale_from_amax(amax)
    for eng, (gxl, gel) in zip(engs, loc):
        eng.bwd_begin(gxl, gel, s)
    for k in range(K - 1, -1, -1):
        for eng in engs:
            eng.bwd_a(k)
        halo.rows_backward_all([eng.dP[eng.part.num_owned:, H:] for eng in engs],
                               [eng.dP[:eng.part.num_owned, H:] for eng in engs])
        for eng in engs:
            eng.bwd_b(k)
    gx0, ge0 = torch.empty_like(x), torch.empty_like(e)
    for p, eng in zip(parts, engs):
        a, b = eng.input_grads()
        gx0[p.node_begin:p.node_begin + p.num_owned] = a
        ge0.index_copy_(0, p.edge_ids, b)
    res.update(g_x=gx0, g_e=ge0, grads=torch.stack([eng.grads for eng in engs]).sum(0))
    return res
