"""ORACLE / TEST INFRASTRUCTURE ONLY -- numpy/scipy restatement of the reference's
one-time integer graph set-up.

Nothing in the product path (graph_physics_b200/) imports this module; it is the
checker used by tests/, by `__graft_entry__.smoke()` and by `bench.py`'s CPU leg.

Parity status
-------------
* Integer side: PINNED.  `tests/test_oracle_golden.py` checks these functions
  against the reference's own golden counts (11,070 one-hop / 32,638 two-hop
  edges on tests/mock_vtu/cylinder_0.vtu, reference
  tests/graphphysics/dataset/test_xdmfdataset.py:171-175,226-231) and the exact
  known-answer tensors of reference tests/graphphysics/utils/test_meshmask.py:59-169,
  and against outputs of the reference's OWN functions (`compute_k_hop_edge_index`,
  `filter_edges`, `meshdata_to_graph` face expansion) executed in the build
  container by tests/golden/make_golden.py.
* The arithmetic that lives in torch_geometric 2.6.1 (requirements.txt:7; not
  vendored under /root/reference, not installed here) is restated from its
  published behaviour: FaceToEdge, to_undirected/coalesce, Cartesian, Distance,
  add_random_edge(force_undirected=True).

Each function cites the reference file:line it follows.  All index outputs are
int64, `edge_index` is `[2, E]` with row 0 = sender, row 1 = receiver.
"""
from __future__ import annotations

import numpy as np


# --------------------------------------------------------------------------- S1
def cells_to_faces(cells: np.ndarray) -> np.ndarray:
    """cells [C, 3|4] -> face [3, F].

    Reference: graphphysics/utils/torch_graph.py:170-186 (same expansion in
    graphphysics/dataset/preprocessing.py:26-46): triangles pass through as
    `cells.T`; a tetrahedron (c0,c1,c2,c3) becomes the four triangles
    (c0,c1,c2), (c1,c2,c3), (c2,c3,c0), (c3,c0,c1), concatenated block-wise
    along the face axis (all first triangles, then all second ones, ...).
    """
    c = np.asarray(cells).T.astype(np.int64)
    if c.shape[0] == 3:
        return c
    if c.shape[0] == 4:
        return np.concatenate(
            [c[0:3], c[1:4], np.stack([c[2], c[3], c[0]]), np.stack([c[3], c[0], c[1]])],
            axis=1,
        )
    raise ValueError("cells must be triangles or tetrahedra")


# --------------------------------------------------------------------------- S2
def coalesce(edge_index: np.ndarray, num_nodes: int) -> np.ndarray:
    """torch_geometric.utils.coalesce: sort by row*N+col, drop duplicates."""
    key = edge_index[0].astype(np.int64) * np.int64(num_nodes) + edge_index[1].astype(np.int64)
    key = np.unique(key)  # sorted + unique
    return np.stack([key // num_nodes, key % num_nodes]).astype(np.int64)


def to_undirected(edge_index: np.ndarray, num_nodes: int) -> np.ndarray:
    """torch_geometric.utils.to_undirected (call site: preprocessing.py:137):
    append the reversed pairs, then coalesce."""
    row, col = edge_index[0], edge_index[1]
    both = np.stack([np.concatenate([row, col]), np.concatenate([col, row])])
    return coalesce(both, num_nodes)


def face_to_edge(face: np.ndarray, num_nodes: int) -> np.ndarray:
    """torch_geometric.transforms.FaceToEdge(remove_faces=False)
    (call sites: preprocessing.py:410,429): the three row-slices face[:2],
    face[1:], face[::2] concatenated, then to_undirected.  The same three slices
    are restated in-repo by jraphphysics/utils/jax_graph.py:100-108."""
    f = np.asarray(face, dtype=np.int64)
    ei = np.concatenate([f[:2], f[1:], f[::2]], axis=1)
    return to_undirected(ei, num_nodes)


# --------------------------------------------------------------------------- S7
def edge_features(pos: np.ndarray, edge_index: np.ndarray) -> np.ndarray:
    """add_edge_features (preprocessing.py:16-23) = T.Cartesian(norm=False) then
    T.Distance(norm=False): [pos[row]-pos[col], ||pos[col]-pos[row]||_2] (fp32)."""
    pos = np.asarray(pos, dtype=np.float32)
    row, col = edge_index
    cart = pos[row] - pos[col]
    d = pos[col] - pos[row]
    dist = np.sqrt((d.astype(np.float32) ** 2).sum(-1, dtype=np.float32)).astype(np.float32)
    return np.concatenate([cart, dist[:, None]], axis=1).astype(np.float32)


# --------------------------------------------------------------------------- S4
def world_edges(edge_index, world_pos, node_type, num_nodes, radius=0.03,
                obstacle=1, normal=0):
    """add_world_edges (preprocessing.py:92-140): cKDTree.query_pairs(radius) on
    the world positions, keep OBSTACLE<->NORMAL pairs only, concatenate in front
    of the mesh edges, to_undirected.  NodeType.NORMAL=0, OBSTACLE=1
    (graphphysics/utils/nodetype.py:4-15)."""
    from scipy.spatial import cKDTree

    tree = cKDTree(np.asarray(world_pos))
    pairs = tree.query_pairs(radius, output_type="ndarray").T.astype(np.int64)
    if pairs.size == 0:
        pairs = np.zeros((2, 0), dtype=np.int64)
    t = np.asarray(node_type)
    m1 = (t[pairs[0]] == obstacle) & (t[pairs[1]] == normal)
    m2 = (t[pairs[0]] == normal) & (t[pairs[1]] == obstacle)
    added = pairs[:, m1 | m2]
    ei = np.concatenate([added, np.asarray(edge_index, dtype=np.int64)], axis=1)
    return to_undirected(ei, num_nodes)


def world_pos_features(edge_attr, world_pos, edge_index):
    """add_world_pos_features (preprocessing.py:143-174): append
    [wpos[s]-wpos[r], ||.||_2] to edge_attr."""
    wp = np.asarray(world_pos, dtype=np.float32)
    s, r = edge_index
    rel = wp[s] - wp[r]
    nrm = np.sqrt((rel ** 2).sum(-1, dtype=np.float32, keepdims=True)).astype(np.float32)
    return np.concatenate([np.asarray(edge_attr, np.float32), rel, nrm], axis=1)


# --------------------------------------------------------------------------- S3
def k_hop_edge_index(edge_index: np.ndarray, num_hops: int, num_nodes: int) -> np.ndarray:
    """compute_k_hop_edge_index (graphphysics/utils/torch_graph.py:16-53):
    A_k <- coalesce(A_k + A_k @ A), self-loops removed after every product,
    (num_hops - 1) times; only the sparsity pattern is returned, in coalesced
    COO order (ascending row, then col)."""
    import scipy.sparse as sp

    ei = coalesce(np.asarray(edge_index, dtype=np.int64), num_nodes)
    a = sp.csr_matrix((np.ones(ei.shape[1], dtype=np.float64), (ei[0], ei[1])),
                      shape=(num_nodes, num_nodes))
    a.data[:] = 1.0
    ak = a.copy()
    for _ in range(num_hops - 1):
        ak = ak + ak @ a
        ak.setdiag(0)
        ak.eliminate_zeros()
        ak.data[:] = 1.0
    coo = ak.tocoo()
    return coalesce(np.stack([coo.row.astype(np.int64), coo.col.astype(np.int64)]), num_nodes)


# --------------------------------------------------------------------------- S5
def add_random_edges(edge_index: np.ndarray, p: float, num_nodes: int, rng: np.random.Generator):
    """_add_random_edges (graphphysics/dataset/dataset.py:104-137) ->
    torch_geometric.utils.add_random_edge(edge_index, p, force_undirected=True):
    round(E*p)//2 undirected node pairs that are neither self-loops nor existing
    edges are drawn, both directions are appended AFTER the existing edges (no
    re-sort).  Random: only the count (E + 2*(round(E*p)//2)), symmetry and
    disjointness are comparable across implementations."""
    ei = np.asarray(edge_index, dtype=np.int64)
    e = ei.shape[1]
    n_pairs = int(round(e * p)) // 2
    have = set((ei[0] * num_nodes + ei[1]).tolist())
    new = []
    seen = set()
    while len(new) < n_pairs:
        a = int(rng.integers(0, num_nodes))
        b = int(rng.integers(0, num_nodes))
        if a == b:
            continue
        lo, hi = (a, b) if a < b else (b, a)
        k = lo * num_nodes + hi
        if k in have or k in seen:
            continue
        seen.add(k)
        new.append((lo, hi))
    new = np.asarray(new, dtype=np.int64).reshape(-1, 2).T
    # to_undirected of the sampled pairs (sorted), appended after the originals
    add = to_undirected(new, num_nodes) if new.size else np.zeros((2, 0), np.int64)
    return np.concatenate([ei, add], axis=1), add


# --------------------------------------------------------------------------- S6
def filter_edges(edge_index: np.ndarray, node_index: np.ndarray, edge_attr=None):
    """filter_edges (graphphysics/utils/meshmask.py:9-37): mask = full(N,-1);
    mask[node_index] = arange(k); remap both ends; keep edges with both >= 0 in
    the ORIGINAL order.  N = max(edge_index)+1 (maybe_num_nodes)."""
    ei = np.asarray(edge_index, dtype=np.int64)
    num_nodes = int(ei.max()) + 1 if ei.size else 0
    node_index = np.asarray(node_index, dtype=np.int64)
    mask = np.full((num_nodes,), -1, dtype=np.int64)
    mask[node_index] = np.arange(node_index.shape[0], dtype=np.int64)
    s, r = mask[ei[0]], mask[ei[1]]
    keep = (s >= 0) & (r >= 0)
    out_attr = None if edge_attr is None else np.asarray(edge_attr)[keep]
    return np.stack([s[keep], r[keep]]), out_attr, keep


def build_masked_graph(x, pos, edge_index, edge_attr, selected):
    """build_masked_graph (meshmask.py:40-70)."""
    ei, ea, keep = filter_edges(edge_index, selected, edge_attr)
    return dict(x=np.asarray(x)[selected],
                pos=None if pos is None else np.asarray(pos)[selected],
                edge_index=ei, edge_attr=ea), keep


def reconstruct_graph(x_full_shape, latent_x, selected, node_mask_token,
                      edge_attr_encoded=None, latent_edge_attr=None, edges_mask=None,
                      edge_mask_token=None):
    """reconstruct_graph (meshmask.py:73-118): features = 0 + node_mask_token;
    features[selected] = latent.x; edges: encoder(edge_attr) + edge_mask_token;
    [edges_mask] = latent.edge_attr."""
    n, f = x_full_shape
    feats = np.zeros((n, f), dtype=np.float32) + np.asarray(node_mask_token, np.float32).reshape(1, -1)
    feats[selected] = latent_x
    ea = None
    if edge_attr_encoded is not None:
        ea = np.asarray(edge_attr_encoded, np.float32) + np.asarray(edge_mask_token, np.float32).reshape(1, -1)
        ea[edges_mask] = latent_edge_attr
    return feats, ea


def masked_indexes(num_nodes: int, masking_ratio: float, perm: np.ndarray) -> np.ndarray:
    """get_masked_indexes (torch_graph.py:285-302): randperm(n)[:int((1-ratio)*n)];
    the permutation itself is RNG-dependent so it is an input here."""
    return np.asarray(perm)[: int((1.0 - masking_ratio) * num_nodes)]


# ------------------------------------------------------- CSR (what the kernels use)
def receiver_csr(edge_index: np.ndarray, num_nodes: int):
    """Receiver-sorted CSR of `edge_index` (stable: ties keep the caller's edge
    order).  This is not a reference function -- it is the layout the reference's
    scatter-sum at edge_index[1] (layers.py:694-696) is replaced with -- but the
    oracle defines it so that the CUDA build can be checked bit-exactly.

    Returns rowptr [N+1] int32, perm [E] int32 (perm[k] = caller edge id of the
    k-th edge in receiver order), senders [E] int32, receivers [E] int32 (both in
    receiver order)."""
    ei = np.asarray(edge_index, dtype=np.int64)
    perm = np.argsort(ei[1], kind="stable").astype(np.int32)
    recv = ei[1][perm]
    send = ei[0][perm]
    counts = np.bincount(recv, minlength=num_nodes)
    rowptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.cumsum(counts, out=rowptr[1:])
    return rowptr.astype(np.int32), perm, send.astype(np.int32), recv.astype(np.int32)
