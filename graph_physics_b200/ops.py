"""Thin typed wrappers over the step kernels of libmgn_b200.so (no autograd here; the
autograd glue lives in processor.py).  All tensors are CUDA, contiguous, fp32 unless
stated; H = 128."""
from __future__ import annotations

import os
from typing import Optional, Sequence

import torch

from . import _lib as L

H = 128


def packed_block_bytes() -> int:
    lib = L.lib()
    lib.mgn_packed_block_bytes.restype = L.ctypes.c_int64
    return int(lib.mgn_packed_block_bytes())


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        raise ValueError("libmgn_b200 takes fp32 tensors, got " + str(t.dtype))
    return t.contiguous()


def pack_block(edge_w: Sequence[torch.Tensor], node_w: Sequence[torch.Tensor],
               out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """edge_w = edge_block.{0,2,4,6}.weight, node_w = node_block.{0,2,4,6}.weight ->
    fp16 tile pack (uint8 [packed_block_bytes])."""
    L.require_device(edge_w[0])
    if tuple(edge_w[0].shape) != (H, 3 * H) or tuple(node_w[0].shape) != (H, 2 * H):
        raise ValueError("libmgn_b200 supports hidden_size=128 with 3H/2H first layers only")
    for wt in list(edge_w[1:]) + list(node_w[1:]):
        if tuple(wt.shape) != (H, H):
            raise ValueError("libmgn_b200 supports hidden_size=128 only")
    if out is None:
        out = torch.empty(packed_block_bytes(), dtype=torch.uint8, device=edge_w[0].device)
    ew = [_f32c(w.detach()) for w in edge_w]
    nw = [_f32c(w.detach()) for w in node_w]
    L.check(L.lib().mgn_pack_block(*[L.p(w) for w in ew], *[L.p(w) for w in nw], L.p(out), L.stream()))
    return out


def pack_blocks(blocks: Sequence[Sequence[torch.Tensor]], out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All blocks of a processor in ONE launch.  blocks[k] = the 8 weights of block k (edge_block.{0,2,4,6}.weight,
    node_block.{0,2,4,6}.weight) -> uint8 [K, packed_block_bytes]."""
    K = len(blocks)
    L.require_device(blocks[0][0])
    keep = []
    for ws in blocks:
        if tuple(ws[0].shape) != (H, 3 * H) or tuple(ws[4].shape) != (H, 2 * H) or \
                any(tuple(w.shape) != (H, H) for w in list(ws[1:4]) + list(ws[5:8])):
            raise ValueError("libmgn_b200 packs hidden_size=128 blocks (3H / 2H first layers) only")
        keep.append([_f32c(w.detach()) for w in ws])
    if out is None:
        out = torch.empty((K, packed_block_bytes()), dtype=torch.uint8, device=blocks[0][0].device)
    arr = (L.ctypes.c_void_p * (8 * K))(*[w.data_ptr() for ws in keep for w in ws])
    L.check(L.lib().mgn_pack_blocks(L.i32(K), arr, L.p(out), L.stream()))
    return out


def p_fwd(x: torch.Tensor, packed: torch.Tensor, b1_edge: torch.Tensor,
          out: Optional[torch.Tensor] = None, chain: Optional[torch.Tensor] = None) -> torch.Tensor:
    """chain: the `chain_state(N)` the node_fwd call that produced `x` (earlier on the same stream) was given, or None."""
    N = x.shape[0]
    if out is None:
        out = torch.empty((N, 2 * H), dtype=torch.float16, device=x.device)
    sched = chain[(N + 127) // 128:] if chain is not None else None
    L.check(L.lib().mgn_p_fwd_c(L.i64(N), L.p(x), L.p(packed), L.p(b1_edge), L.p(out), L.p(chain), L.p(sched),
                                L.stream()))
    return out


# The sum at the receivers is deterministic and atomics-free (plain stores + a fix-up pass over the slot-boundary
# carries, csrc/edge_fwd_hl.cu aggr_fixup_kernel).
DETERMINISTIC = True


def carry_buffer(num_edges: int, slot_rows: int, device) -> torch.Tensor:
    """Workspace of the deterministic aggregation: one 128-float row per `slot_rows`-edge slot (8 for edge_fwd, 32 for
    edge_fwd_hl).  Contents need no initialisation: only rows a kernel wrote are read back."""
    return torch.empty(((num_edges + slot_rows - 1) // slot_rows, H), dtype=torch.float32, device=device)


def edge_fwd(e_in: torch.Tensor, rowmap: Optional[torch.Tensor], P: torch.Tensor, senders: torch.Tensor,
             receivers: torch.Tensor, num_nodes: int, packed: torch.Tensor, b2, b3, b4, scale,
             e_out: Optional[torch.Tensor] = None, aggr: Optional[torch.Tensor] = None, saves=None,
             e16_out: Optional[torch.Tensor] = None, e16_in: Optional[torch.Tensor] = None,
             e16_next: Optional[torch.Tensor] = None, out_in_csr_order: bool = False,
             e16lo_next: Optional[torch.Tensor] = None, write_rows: bool = True,
             carry: Optional[torch.Tensor] = None, norm_dim: int = H, rowptr: Optional[torch.Tensor] = None):
    """saves: optional 3 fp16 [E,128] buffers receiving the hidden activations (row = CSR position;
    used by the parity tests only -- the backward kernel recomputes them).
    e16_out: optional `e16_image(E)` buffer receiving the fp16 tile images of e_in that edge_bwd bulk-loads.
    e16_in: the tile images of e_in if a previous step already made them (its e16_next): the operand tiles then
    arrive by bulk copy and e_in is read for the residual only.  e16_next: receives the images of e_out.
    out_in_csr_order: with a rowmap, gather e_in through it but write e_out in CSR order (fuses the entry permutation
    of the processor loop into its first step).
    e16lo_next: receives the images of e_out - fp16(e_out) (the remainder stream edge_fwd_hl reads); with
    write_rows=False the fp32 rows of e_out are then not written (returns None for them)."""
    E = e_in.shape[0]
    s1, s2, s3 = saves if saves is not None else (None, None, None)
    if not write_rows:
        if e16_next is None or e16lo_next is None:
            raise ValueError("write_rows=False needs e16_next and e16lo_next")
        e_out = None
    elif e_out is None:
        e_out = torch.empty_like(e_in)
    if aggr is None:
        aggr = torch.empty((num_nodes, H), dtype=torch.float32, device=e_in.device)
    if carry is None and DETERMINISTIC and E > 0:
        carry = carry_buffer(E, 8, e_in.device)
    L.check(L.lib().mgn_edge_fwd2(L.i64(E), L.i64(num_nodes), L.p(e_in), L.p(rowmap), L.p(e_out), L.p(P),
                                  L.p(senders), L.p(receivers), L.p(aggr), L.p(packed), L.p(b2), L.p(b3), L.p(b4),
                                  L.p(scale), L.p(s1), L.p(s2), L.p(s3), L.p(e16_out), L.p(e16_in), L.p(e16_next),
                                  L.p(e16lo_next), L.i32(1 if out_in_csr_order else 0),
                                  L.p(carry), L.p(rowptr), L.i32(norm_dim), L.stream()))
    return e_out, aggr


# The processor loop keeps the edge stream of steps 1..K-1 as two fp16 tile-image streams (csrc/edge_fwd_hl.cu);
# MGN_EDGE_FWD_HL=0 falls back to the fp32-row kernel for every step.
EDGE_FWD_HL = bool(int(__import__("os").environ.get("MGN_EDGE_FWD_HL", "1")))


def edge_fwd_hl(hi_in: torch.Tensor, lo_in: torch.Tensor, P: torch.Tensor, senders: torch.Tensor,
                receivers: torch.Tensor, num_nodes: int, packed: torch.Tensor, b2, b3, b4, scale,
                hi_out: Optional[torch.Tensor] = None, lo_out: Optional[torch.Tensor] = None,
                e_out: Optional[torch.Tensor] = None, aggr: Optional[torch.Tensor] = None,
                carry: Optional[torch.Tensor] = None, norm_dim: int = H, rowptr: Optional[torch.Tensor] = None,
                h_saves=None):
    """h_saves: optional 3 `e16_image(E)` buffers receiving the edge MLP's hidden activations (training: what
    `edge_bwd_saved` starts from instead of recomputing them).
    Edge update of one step (layers.py:687-698) on the image + remainder form of the edge stream:
    e_in = hi_in + lo_in (`e16_image` buffers, CSR tile order).  hi_out / lo_out receive the same form of e_out,
    e_out (optional, [E,128] fp32, CSR order) its fp32 rows.  -> aggr"""
    E = senders.shape[0]
    if aggr is None:
        aggr = torch.empty((num_nodes, H), dtype=torch.float32, device=hi_in.device)
    if carry is None and DETERMINISTIC and E > 0:
        carry = carry_buffer(E, 32, hi_in.device)
    hs = h_saves if h_saves is not None else (None, None, None)
    L.check(L.lib().mgn_edge_fwd_hl(L.i64(E), L.i64(num_nodes), L.p(hi_in), L.p(lo_in), L.p(hi_out), L.p(lo_out),
                                    L.p(e_out), L.p(P), L.p(senders), L.p(receivers), L.p(aggr), L.p(packed),
                                    L.p(b2), L.p(b3), L.p(b4), L.p(scale), L.p(carry), L.p(rowptr),
                                    L.i32(norm_dim), L.p(hs[0]), L.p(hs[1]), L.p(hs[2]), L.stream()))
    return aggr


def e16_image(num_edges: int, device) -> torch.Tensor:
    """Buffer for the fp16 tile images of an edge-latent version: one 32 KB swizzled 128 x 128 tile per 128 edges."""
    tiles = (num_edges + 127) // 128
    return torch.empty((tiles, 128 * H), dtype=torch.float16, device=device)


def node_saves(num_rows: int, device):
    """The three buffers node_fwd fills with the node MLP's hidden activations for node_bwd: fp16 tile images (the
    128 x 128 swizzled operand tiles as they sit in shared memory, bulk-stored / bulk-loaded)."""
    return tuple(e16_image(num_rows, device) for _ in range(3))


def image_rows(img: torch.Tensor, num_rows: int) -> torch.Tensor:
    """[num_rows,128] fp32 view of a tile-image buffer (`e16_image` / `node_saves`): undoes the 128-byte swizzle of
    csrc/tc05.cuh tile_off.  For tests and debugging; the kernels never need the rows."""
    t = img.view(-1, 2, 128, 8, 8).float()                       # tile, 64-column panel, row, 16-byte slot, 8 halfs
    rows = torch.arange(128, device=img.device)
    chunk = torch.arange(8, device=img.device)
    slot = chunk[None, :] ^ (rows[:, None] & 7)                  # chunk c of row r sits in slot c ^ (r & 7)
    un = torch.gather(t, 3, slot[None, None, :, :, None].expand(t.shape[0], 2, 128, 8, 8))
    return un.permute(0, 2, 1, 3, 4).reshape(-1, 128)[:num_rows]


def node_fwd(x: torch.Tensor, aggr: torch.Tensor, packed: torch.Tensor, b1, b2, b3, b4, scale,
             x_out: Optional[torch.Tensor] = None, saves=None, norm: bool = True, norm_dim: int = H,
             chain: Optional[torch.Tensor] = None) -> torch.Tensor:
    """saves: `node_saves(N)` buffers (tile images), or None.  norm=False: the block has no RMSNorm (decoder).
    chain: `chain_state(N)` shared with the p_fwd call that consumes x_out next (it MUST follow: the flags this call
    sets are cleared by that consumer)."""
    N = x.shape[0]
    if saves is not None and any(sv.numel() < ((N + 127) // 128) * 128 * H for sv in saves):
        raise ValueError("node_fwd: saves must be node_saves(N) tile-image buffers")
    if x_out is None:
        x_out = torch.empty_like(x)
    s1, s2, s3 = saves if saves is not None else (None, None, None)
    L.check(L.lib().mgn_node_fwd3(L.i64(N), L.p(x), L.p(aggr), L.p(x_out), L.p(packed), L.p(b1), L.p(b2), L.p(b3),
                                  L.p(b4), L.p(scale), L.p(s1), L.p(s2), L.p(s3), L.i32(0 if norm else 1), L.i32(norm_dim),
                                  L.p(chain), L.stream()))
    return x_out


# ------------------------------------------------------------------ backward
# flat gradient block of one GraphNetBlock, in state-dict order (csrc/mlp_fwd_abi.cuh GB_*)
GRAD_LAYOUT = (
    ("edge_block.0.weight", (H, 3 * H)), ("edge_block.0.bias", (H,)),
    ("edge_block.2.weight", (H, H)), ("edge_block.2.bias", (H,)),
    ("edge_block.4.weight", (H, H)), ("edge_block.4.bias", (H,)),
    ("edge_block.6.weight", (H, H)), ("edge_block.6.bias", (H,)),
    ("edge_block.7.scale", (H,)),
    ("node_block.0.weight", (H, 2 * H)), ("node_block.0.bias", (H,)),
    ("node_block.2.weight", (H, H)), ("node_block.2.bias", (H,)),
    ("node_block.4.weight", (H, H)), ("node_block.4.bias", (H,)),
    ("node_block.6.weight", (H, H)), ("node_block.6.bias", (H,)),
    ("node_block.7.scale", (H,)),
)
GRAD_BLOCK_FLOATS = sum(int(torch.Size(s).numel()) for _, s in GRAD_LAYOUT)   # 181,504


def split_grad_block(flat: torch.Tensor):
    """flat [GRAD_BLOCK_FLOATS] -> {name: view} in GRAD_LAYOUT order (no copies)."""
    out, off = {}, 0
    for name, shape in GRAD_LAYOUT:
        n = int(torch.Size(shape).numel())
        out[name] = flat[off:off + n].view(shape)
        off += n
    return out


def edge_bwd(e_in, rowmap, g_eout, g_aggr, P, senders, receivers, num_nodes, packed, b2, b3, b4, scale,
             grad_block, g_e=None, dz16=None, e16=None, inplace=False, norm_dim: int = H):
    """-> (g_e [E,128], dz16: fp16 tile images of dz1 for p_bwd).  g_eout may be None (zero upstream edge gradient).
    `inplace` writes g_e over g_eout (each tile is read before it is written).
    e16: the forward kernel's fp16 tile images of e_in (then e_in may be None)."""
    E = senders.shape[0]
    dev = g_aggr.device
    if g_e is None:
        g_e = g_eout if (inplace and g_eout is not None) else torch.empty((E, H), dtype=torch.float32, device=dev)
    g_src = g_eout
    if dz16 is None:
        dz16 = e16_image(E, dev)
    L.check(L.lib().mgn_edge_bwd(L.i64(E), L.i64(num_nodes), L.p(e_in), L.p(e16), L.p(rowmap), L.p(g_src),
                                 L.p(g_aggr), L.p(P), L.p(senders), L.p(receivers), L.p(g_e), L.p(dz16), L.p(packed),
                                 L.p(b2), L.p(b3), L.p(b4), L.p(scale), L.p(grad_block), L.i32(norm_dim), L.stream()))
    return g_e, dz16


# MGN_SAVE_EDGE_HIDDEN=0: the processor loop does not keep the edge MLP's hidden activations (768 B per edge and step)
# and its backward recomputes them (mgn_edge_bwd) -- smaller memory footprint, ~35 % slower edge backward.
SAVE_EDGE_HIDDEN = bool(int(__import__("os").environ.get("MGN_SAVE_EDGE_HIDDEN", "1")))


def edge_bwd_saved(h_saves, g_e, g_aggr, receivers, num_nodes, packed, b4, scale, grad_block, dz16=None,
                   norm_dim: int = H):
    """Edge backward of a processor-loop step from the hidden activations `edge_fwd_hl(h_saves=...)` kept.  g_e [E,128]
    (CSR order) is updated IN PLACE from the upstream edge gradient to the gradient of the step's input edges.
    -> (g_e, dz16)"""
    E = receivers.shape[0]
    if dz16 is None:
        dz16 = e16_image(E, g_e.device)
    h1, h2, h3 = h_saves
    L.check(L.lib().mgn_edge_bwd_saved(L.i64(E), L.i64(num_nodes), L.p(h1), L.p(h2), L.p(h3), L.p(g_aggr),
                                       L.p(receivers), L.p(g_e), L.p(dz16), L.p(packed), L.p(b4), L.p(scale),
                                       L.p(grad_block), L.i32(norm_dim), L.stream()))
    return g_e, dz16


# Tile-level chaining of two launches (csrc/mlp_common.cuh): the producer publishes ready[tile] after a tile's last
# store, the consumer -- whose CTAs start on the SMs the producer's wave tail leaves idle -- takes tiles from a counter
# and waits for the flag.  MGN_CHAIN=0 switches it off (plain stream order).
CHAIN = os.environ.get("MGN_CHAIN", "1") != "0"


def chain_state(num_rows: int, device) -> torch.Tensor:
    """int32 [tiles + 2]: ready flags of one producer / consumer pair + {next tile, CTAs done}; starts (and is left) at
    zero.  ready = state[:tiles], sched = state[tiles:]."""
    return torch.zeros((num_rows + 127) // 128 + 2, dtype=torch.int32, device=device)


def node_bwd(g_xout, saves, packed, b4, scale, grad_block, g_x=None, g_aggr=None, dz1=None, norm: bool = True,
             norm_dim: int = H, chain: Optional[torch.Tensor] = None, chain_in: Optional[torch.Tensor] = None):
    """-> (g_x direct part, g_aggr, dz1 fp16).  chain: `chain_state(N)` shared with the node_wgrad1 call that follows.
    chain_in: the `chain_state(N)` the p_bwd call that wrote `g_xout` (earlier on the same stream) was given."""
    N = g_xout.shape[0]
    if g_x is None:
        g_x = torch.empty_like(g_xout)
    if g_aggr is None:
        g_aggr = torch.empty_like(g_xout)
    if dz1 is None:
        dz1 = torch.empty((N, H), dtype=torch.float16, device=g_xout.device)
    h1, h2, h3 = saves
    L.check(L.lib().mgn_node_bwd3(L.i64(N), L.p(g_xout), L.p(h1), L.p(h2), L.p(h3), L.p(g_x), L.p(g_aggr), L.p(dz1),
                                  L.p(packed), L.p(b4), L.p(scale), L.p(grad_block), L.i32(0 if norm else 1), L.i32(norm_dim),
                                  L.p(chain_in), L.p(chain), L.stream()))
    return g_x, g_aggr, dz1


def edge_wgrad1(dz16, e16, num_edges, grad_block):
    """dWe += dz1^T e (edge_block.0.weight[:, :128]) from the tile images of dz1 (edge_bwd) and of e (edge_fwd)."""
    L.check(L.lib().mgn_edge_wgrad1(L.i64(num_edges), L.p(dz16), L.p(e16), L.p(grad_block), L.stream()))


def dp_gather(dz16, csr, dP=None):
    """dP [N,256] fp16 = [sum of dz1 over received edges | over sent edges], gathered from edge_bwd's tile images in
    a fixed order (no atomics).  csr: graph.CSRGraph."""
    rowptr_s, out_pos = csr.by_sender
    if dP is None:
        dP = torch.empty((csr.num_nodes, 2 * H), dtype=torch.float16, device=dz16.device)
    L.check(L.lib().mgn_dp_gather(L.i64(csr.num_nodes), L.p(dz16), L.p(csr.rowptr), L.p(rowptr_s), L.p(out_pos),
                                  L.p(dP), L.stream()))
    return dP


def p_bwd(dP, x, g_x, packed, grad_block, chain: Optional[torch.Tensor] = None):
    """g_x += dP [W_i ; W_j] (in place); dW_i, dW_j, db1 into grad_block.  chain: `chain_state(N)` shared with the
    node_bwd call that consumes g_x next (it MUST follow: the flags this call sets are cleared by that consumer)."""
    L.check(L.lib().mgn_p_bwd_c(L.i64(x.shape[0]), L.p(dP), L.p(x), L.p(g_x), L.p(packed), L.p(grad_block),
                                L.p(chain), L.stream()))
    return g_x


def node_wgrad1(dz1, x, aggr, grad_block, chain: Optional[torch.Tensor] = None):
    """chain: the `chain_state(N)` the node_bwd call JUST BEFORE this one (same stream) was given, or None."""
    N = x.shape[0]
    sched = chain[(N + 127) // 128:] if chain is not None else None
    L.check(L.lib().mgn_node_wgrad1c(L.i64(N), L.p(dz1), L.p(x), L.p(aggr), L.p(grad_block), L.p(chain), L.p(sched),
                                     L.stream()))
