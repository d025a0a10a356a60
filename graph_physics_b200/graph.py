"""Device-side graph set-up with the reference's function names and results
(SURVEY.md 8(a) S1-S7).  Every function takes CUDA tensors and calls
libmgn_b200.so; index results are bit-exact with the reference's
torch_geometric / ATen path (int64 `edge_index` [2,E], ascending (row, col)).

reference function                                   | here
-----------------------------------------------------|---------------------------
meshdata_to_graph cells->face (torch_graph.py:170-186) | cells_to_faces
T.FaceToEdge + to_undirected (preprocessing.py:410,429,137) | face_to_edge, to_undirected
compute_k_hop_edge_index (torch_graph.py:16-53)      | compute_k_hop_edge_index
filter_edges / build_masked_graph / reconstruct_graph (meshmask.py:9-118) | same names
add_edge_features = Cartesian+Distance (preprocessing.py:16-23) | edge_features
add_world_pos_features (preprocessing.py:143-174)    | world_pos_features
add_world_edges (preprocessing.py:92-140)            | world_pairs, add_world_edges
_add_random_edges (dataset/dataset.py:104-137)       | add_random_edges
PyG Batch collation of identical meshes (train.py:164-165) | tile_batch
(new) receiver-sorted CSR replacing the scatter-sum  | build_csr -> CSRGraph
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import _lib as L


def _i64c(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.int64).contiguous()


def _ws(nbytes: int, device) -> torch.Tensor:
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


def cells_to_faces(cells: torch.Tensor) -> torch.Tensor:
    """cells [C,3] or [C,4] (rows = cells, as meshio gives them) -> face [3,F] int64."""
    L.require_device(cells)
    c = _i64c(cells.t())
    if c.shape[0] == 3:
        return c
    if c.shape[0] != 4:
        raise ValueError("cells must be triangles or tetrahedra")
    T = c.shape[1]
    face = torch.empty((3, 4 * T), dtype=torch.int64, device=c.device)
    L.check(L.lib().mgn_tets_to_faces(L.p(c), L.i64(T), L.p(face), L.stream()))
    return face


def _sort_unique_to_edges(keys: torch.Tensor, num_nodes: int) -> torch.Tensor:
    n = keys.numel()
    dev = keys.device
    if n == 0:
        return torch.empty((2, 0), dtype=torch.int64, device=dev)
    out = torch.empty_like(keys)
    wsb = L.lib().mgn_sort_unique_workspace_bytes(L.i64(n))
    ws = _ws(wsb, dev)
    cnt = L.ctypes.c_int64(0)
    L.check(L.lib().mgn_sort_unique_keys(L.p(keys), L.i64(n), L.i64(num_nodes), L.p(out), L.p(ws),
                                         L.i64(wsb), L.ctypes.byref(cnt), L.stream()))
    E = int(cnt.value)
    ei = torch.empty((2, E), dtype=torch.int64, device=dev)
    L.check(L.lib().mgn_keys_to_edge_index(L.p(out), L.i64(E), L.i64(num_nodes), L.p(ei), L.stream()))
    return ei


def face_to_edge(face: torch.Tensor, num_nodes: int) -> torch.Tensor:
    """T.FaceToEdge(remove_faces=False): unique undirected edges of all triangle sides."""
    L.require_device(face)
    f = _i64c(face)
    F = f.shape[1]
    keys = torch.empty(6 * F, dtype=torch.int64, device=f.device)
    L.check(L.lib().mgn_face_keys(L.p(f), L.i64(F), L.i64(num_nodes), L.p(keys), L.stream()))
    return _sort_unique_to_edges(keys, num_nodes)


def _edge_keys(edge_index: torch.Tensor, num_nodes: int, both: bool) -> torch.Tensor:
    ei = _i64c(edge_index)
    E = ei.shape[1]
    keys = torch.empty(E * (2 if both else 1), dtype=torch.int64, device=ei.device)
    L.check(L.lib().mgn_edge_keys(L.p(ei), L.i64(E), L.i64(num_nodes), L.i32(1 if both else 0), L.p(keys),
                                  L.stream()))
    return keys


def to_undirected(edge_index: torch.Tensor, num_nodes: int) -> torch.Tensor:
    L.require_device(edge_index)
    return _sort_unique_to_edges(_edge_keys(edge_index, num_nodes, True), num_nodes)


def coalesce(edge_index: torch.Tensor, num_nodes: int) -> torch.Tensor:
    L.require_device(edge_index)
    return _sort_unique_to_edges(_edge_keys(edge_index, num_nodes, False), num_nodes)


def compute_k_hop_edge_index(edge_index: torch.Tensor, num_hops: int, num_nodes: int) -> torch.Tensor:
    """A_k <- pattern(A_k + A_k A) without self-loops, (num_hops-1) times; coalesced order."""
    L.require_device(edge_index)
    a = coalesce(edge_index, num_nodes)
    ak = a
    lib = L.lib()
    for _ in range(num_hops - 1):
        E, Ek = a.shape[1], ak.shape[1]
        if E == 0 or Ek == 0:
            break
        wsb = lib.mgn_khop_workspace_bytes(L.i64(E), L.i64(Ek), L.i64(num_nodes))
        ws = _ws(wsb, a.device)
        tot = L.ctypes.c_int64(0)
        L.check(lib.mgn_khop_count(L.p(a), L.i64(E), L.p(ak), L.i64(Ek), L.i64(num_nodes), L.p(ws), L.i64(wsb),
                                   L.ctypes.byref(tot), L.stream()))
        keys = torch.empty(int(tot.value), dtype=torch.int64, device=a.device)
        L.check(lib.mgn_khop_expand(L.p(a), L.i64(E), L.p(ak), L.i64(Ek), L.i64(num_nodes), L.p(ws), L.p(keys),
                                    L.stream()))
        ak = _sort_unique_to_edges(keys, num_nodes)
    return ak


def edge_features(pos: torch.Tensor, edge_index: torch.Tensor) -> torch.Tensor:
    """[pos[row]-pos[col], ||pos[col]-pos[row]||] fp32, width D+1."""
    L.require_device(pos)
    pos = pos.to(torch.float32).contiguous()
    ei = _i64c(edge_index)
    E, D = ei.shape[1], pos.shape[1]
    out = torch.empty((E, D + 1), dtype=torch.float32, device=pos.device)
    L.check(L.lib().mgn_edge_features(L.p(pos), L.i32(D), L.p(ei), L.i64(E), L.p(out), L.stream()))
    return out


def world_pos_features(edge_attr: torch.Tensor, world_pos: torch.Tensor, edge_index: torch.Tensor) -> torch.Tensor:
    """append [wpos[s]-wpos[r], ||.||] (same arithmetic as edge_features on world_pos)."""
    return torch.cat([edge_attr, edge_features(world_pos, edge_index)], dim=-1)


def world_pairs(world_pos: torch.Tensor, node_type: torch.Tensor, radius: float = 0.03, obstacle: int = 1,
                normal: int = 0, sample: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All (OBSTACLE node, NORMAL node) pairs at world distance <= radius: [2,M] int64, row 0 = obstacle.  The
    pairs cKDTree.query_pairs + the two node-type masks keep in add_world_edges (preprocessing.py:103-131), found
    with a uniform cell list on the device.  `sample` [N] restricts pairs to nodes of the same sample (the
    reference runs per sample, before collation).  NodeType.NORMAL = 0, OBSTACLE = 1 (utils/nodetype.py:4-15)."""
    import ctypes
    L.require_device(world_pos)
    pos = world_pos.to(torch.float32).contiguous()
    N, D = pos.shape
    dev = pos.device
    if D not in (2, 3):
        raise ValueError("world positions must be 2-D or 3-D")
    if not radius > 0:
        raise ValueError("radius must be positive")
    if N == 0:
        return torch.empty((2, 0), dtype=torch.int64, device=dev)
    t = node_type.reshape(-1)
    cls = ((t == obstacle).to(torch.uint8) + 2 * (t == normal).to(torch.uint8)).contiguous()
    smp = sample.to(torch.int32).contiguous() if sample is not None else None
    lo = pos.min(0).values.double().cpu()
    hi = pos.max(0).values.double().cpu()
    cell = 1.001 * float(radius)
    dims = [int((float(hi[a]) - float(lo[a])) // cell) + 1 if a < D else 1 for a in range(3)]
    nsmp = int(smp.max()) + 1 if smp is not None else 1
    if nsmp * dims[0] * dims[1] * dims[2] >= 2 ** 62:
        raise ValueError("world-edge cell grid too large for the radius")
    origin = (ctypes.c_double * 3)(*[float(lo[a]) if a < D else 0.0 for a in range(3)])
    cdims = (ctypes.c_int64 * 3)(*dims)
    keys = torch.empty(N, dtype=torch.int64, device=dev)
    L.check(L.lib().mgn_world_cell_keys(L.p(pos), L.i32(D), L.i64(N), L.p(smp), origin, ctypes.c_double(radius),
                                        cdims, L.p(keys), L.stream()))
    keys_sorted, order = torch.sort(keys)
    order = order.to(torch.int32)
    counts = torch.empty(N, dtype=torch.int64, device=dev)
    args = (L.p(pos), L.i32(D), L.i64(N), L.p(cls), L.p(smp), L.p(keys_sorted), L.p(order), origin,
            ctypes.c_double(radius), cdims)
    L.check(L.lib().mgn_world_pairs(*args, L.p(counts), L.p(None), L.p(None), L.i64(0), L.stream()))
    ends = torch.cumsum(counts, 0)
    M = int(ends[-1])
    pairs = torch.empty((2, M), dtype=torch.int64, device=dev)
    if M:
        offs = (ends - counts).contiguous()
        L.check(L.lib().mgn_world_pairs(*args, L.p(None), L.p(offs), L.p(pairs), L.i64(M), L.stream()))
    return pairs


def add_world_edges(edge_index: torch.Tensor, world_pos: torch.Tensor, node_type: torch.Tensor, num_nodes: int,
                    radius: float = 0.03, obstacle: int = 1, normal: int = 0,
                    sample: Optional[torch.Tensor] = None) -> torch.Tensor:
    """add_world_edges (preprocessing.py:92-140) on tensors: world pairs in front of the mesh edges, to_undirected."""
    pairs = world_pairs(world_pos, node_type, radius, obstacle, normal, sample)
    return to_undirected(torch.cat([pairs, _i64c(edge_index)], dim=1), num_nodes)


def add_random_edges(edge_index: torch.Tensor, p: float, num_nodes: int,
                     generator: Optional[torch.Generator] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """_add_random_edges (dataset/dataset.py:104-137) -> PyG add_random_edge(edge_index, p, force_undirected=True):
    round(E p) // 2 new undirected pairs (no self loops, none already present, no duplicates), both directions
    appended AFTER the existing edges.  Returns (edge_index', added [2, 2*pairs]).  Random: comparable with the
    reference by count, symmetry and disjointness only."""
    L.require_device(edge_index)
    ei = _i64c(edge_index)
    dev = ei.device
    E, N = ei.shape[1], int(num_nodes)
    want = int(round(E * p)) // 2
    if want == 0 or N < 2:
        return ei, torch.empty((2, 0), dtype=torch.int64, device=dev)
    lo, hi = torch.minimum(ei[0], ei[1]), torch.maximum(ei[0], ei[1])
    have = torch.unique(lo * N + hi)
    free = N * (N - 1) // 2 - int(((have // N) != (have % N)).sum())
    if want > free:
        raise ValueError("not enough absent node pairs for the requested random edges")
    new = torch.empty(0, dtype=torch.int64, device=dev)
    while new.numel() < want:
        m = 2 * (want - int(new.numel())) + 64
        a = torch.randint(0, N, (m,), device=dev, generator=generator)
        b = torch.randint(0, N, (m,), device=dev, generator=generator)
        k = torch.minimum(a, b) * N + torch.maximum(a, b)
        k = k[a != b]
        pos = torch.searchsorted(have, k).clamp_(max=have.numel() - 1)
        k = k[have[pos] != k] if have.numel() else k
        new = torch.unique(torch.cat([new, k]))
        if new.numel() > want:      # unique() sorted them: drop a random subset, not the largest keys
            keep = torch.randperm(new.numel(), device=dev, generator=generator)[:want]
            new = new[keep]
    pairs = torch.stack([new // N, new % N])
    added = to_undirected(pairs, N)
    return torch.cat([ei, added], dim=1), added


def filter_edges(edge_index: torch.Tensor, node_index: torch.Tensor, edge_attr: Optional[torch.Tensor] = None):
    """Returns (edge_index', edge_attr' or None, edges_mask bool [E]); order preserved."""
    L.require_device(edge_index)
    ei = _i64c(edge_index)
    ni = _i64c(node_index.to(ei.device))
    E = ei.shape[1]
    N = int(ei.max().item()) + 1 if E > 0 else 0     # maybe_num_nodes
    dev = ei.device
    out = torch.empty((2, E), dtype=torch.int64, device=dev)
    kept = torch.empty(E, dtype=torch.int64, device=dev)
    mask = torch.zeros(E, dtype=torch.uint8, device=dev)
    cnt = L.ctypes.c_int64(0)
    if E > 0:
        wsb = L.lib().mgn_filter_workspace_bytes(L.i64(E), L.i64(N))
        ws = _ws(wsb, dev)
        L.check(L.lib().mgn_filter_edges(L.p(ei), L.i64(E), L.p(ni), L.i64(ni.numel()), L.i64(N), L.p(out),
                                         L.p(kept), L.p(mask), L.p(ws), L.i64(wsb), L.ctypes.byref(cnt),
                                         L.stream()))
    k = int(cnt.value)
    new_ei = out[:, :k].contiguous()
    new_attr = None if edge_attr is None else edge_attr.index_select(0, kept[:k])
    return new_ei, new_attr, mask.bool()


def build_masked_graph(masked_graph, selected_indexes: torch.Tensor):
    """meshmask.build_masked_graph: mutates and returns (graph, edges_mask)."""
    ei, ea, edges_mask = filter_edges(masked_graph.edge_index, selected_indexes,
                                      getattr(masked_graph, "edge_attr", None))
    masked_graph.edge_index = ei
    masked_graph.x = masked_graph.x[selected_indexes]
    if getattr(masked_graph, "pos", None) is not None:
        masked_graph.pos = masked_graph.pos[selected_indexes]
    if ea is not None:
        masked_graph.edge_attr = ea
    return masked_graph, edges_mask


def reconstruct_graph(graph, latent_masked_graph, selected_indexes, node_mask_token, edges_mask,
                      edge_encoder=None, edge_mask_token=None):
    """meshmask.reconstruct_graph: [MASK] tokens everywhere, latent rows re-inserted."""
    n, f = graph.x.shape
    features = torch.zeros(n, f, device=graph.x.device) + node_mask_token.expand(n, -1)
    features[selected_indexes] = latent_masked_graph.x
    latent = graph.clone()
    latent.x = features
    if getattr(graph, "edge_attr", None) is not None:
        ne = graph.edge_attr.shape[0]
        ea = edge_encoder(latent.edge_attr) + edge_mask_token.expand(ne, -1)
        ea[edges_mask] = latent_masked_graph.edge_attr
        latent.edge_attr = ea
    return latent


def tile_batch(edge_index: torch.Tensor, num_nodes: int, copies: int) -> torch.Tensor:
    """Block-diagonal batch of `copies` identical meshes, node offsets as PyG Batch."""
    offs = (torch.arange(copies, device=edge_index.device, dtype=torch.int64) * num_nodes).view(1, copies, 1)
    return (edge_index.view(2, 1, -1) + offs).reshape(2, -1).contiguous()


@dataclass
class CSRGraph:
    """Receiver-sorted CSR of an edge_index (stable in the caller's edge order)."""
    num_nodes: int
    num_edges: int
    rowptr: torch.Tensor      # [N+1] int32
    perm: torch.Tensor        # [E] int32: caller edge id of CSR edge k
    senders: torch.Tensor     # [E] int32, CSR order
    receivers: torch.Tensor   # [E] int32, CSR order (non-decreasing)
    identity: bool            # perm == arange (caller order already receiver-sorted)

    @property
    def perm_long(self) -> torch.Tensor:
        pl = getattr(self, "_perm_long", None)
        if pl is None:
            pl = self.perm.long()
            object.__setattr__(self, "_perm_long", pl)
        return pl

    @property
    def by_sender(self):
        """(rowptr_s [N+1] int32, out_pos [E] int32): for node n, out_pos[rowptr_s[n]:rowptr_s[n+1]] are the
        CSR positions of the edges n SENDS (the gather lists of the deterministic backward of x[row])."""
        bs = getattr(self, "_by_sender", None)
        if bs is None:
            flipped = torch.stack([self.receivers.long(), self.senders.long()])   # sort key = sender
            c2 = build_csr(flipped, self.num_nodes)
            bs = (c2.rowptr, c2.perm)
            object.__setattr__(self, "_by_sender", bs)
        return bs


def build_csr(edge_index: torch.Tensor, num_nodes: int) -> CSRGraph:
    L.require_device(edge_index)
    ei = _i64c(edge_index)
    E = ei.shape[1]
    dev = ei.device
    rowptr = torch.empty(num_nodes + 1, dtype=torch.int32, device=dev)
    perm = torch.empty(E, dtype=torch.int32, device=dev)
    send = torch.empty(E, dtype=torch.int32, device=dev)
    recv = torch.empty(E, dtype=torch.int32, device=dev)
    wsb = L.lib().mgn_csr_workspace_bytes(L.i64(max(E, 1)))
    ws = _ws(wsb, dev)
    L.check(L.lib().mgn_build_csr(L.p(ei), L.i64(E), L.i64(num_nodes), L.p(rowptr), L.p(perm), L.p(send),
                                  L.p(recv), L.p(ws), L.i64(wsb), L.stream()))
    ident = bool(E == 0 or torch.equal(perm, torch.arange(E, dtype=torch.int32, device=dev)))
    return CSRGraph(num_nodes, E, rowptr, perm, send, recv, ident)
