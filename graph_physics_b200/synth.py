"""Synthetic 3D meshes for the large-mesh workloads (SURVEY.md 8(d) "Synthetic inputs": Kuhn-split structured
tetrahedral grids for BASELINE.json configs 3-5).  Benchmarks and tests only; nothing here is on the hot path.

A Kuhn (Freudenthal) split cuts every grid cube into the 6 tetrahedra that share the cube's main diagonal: for every
order (a, b, c) of the three axes, the tetrahedron {v, v+a, v+a+b, v+a+b+c}.  Nodes are numbered x-major,
id = (ix * ny + iy) * nz + iz.
"""
from __future__ import annotations

import itertools

import torch

# the 7 undirected edge directions of the Kuhn split: 3 axes, 3 face diagonals, the body diagonal
KUHN_DIRECTIONS = ((1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (0, 1, 1), (1, 0, 1), (1, 1, 1))


def kuhn_grid_cells(nx: int, ny: int, nz: int, device="cpu") -> torch.Tensor:
    """cells [6 (nx-1)(ny-1)(nz-1), 4] int64 (rows = tetrahedra, as meshio gives them)."""
    ix, iy, iz = torch.meshgrid(torch.arange(nx - 1, device=device), torch.arange(ny - 1, device=device),
                                torch.arange(nz - 1, device=device), indexing="ij")
    base = torch.stack([ix.flatten(), iy.flatten(), iz.flatten()], dim=1)          # [C,3]
    unit = torch.eye(3, dtype=torch.int64, device=device)
    stride = torch.tensor([ny * nz, nz, 1], dtype=torch.int64, device=device)
    tets = []
    for order in itertools.permutations(range(3)):
        v, corners = base, [base]
        for ax in order:
            v = v + unit[ax]
            corners.append(v)
        tets.append(torch.stack([(c * stride).sum(1) for c in corners], dim=1))     # [C,4]
    return torch.cat(tets, dim=0)


def kuhn_grid_edges(nx: int, ny: int, nz: int, device="cpu") -> torch.Tensor:
    """edge_index [2,E] int64 of the Kuhn grid, both directions, sorted by sender then receiver: exactly what
    face_to_edge(cells_to_faces(kuhn_grid_cells(...))) returns (the reference's T.FaceToEdge order), built from the 7
    edge directions without materialising cells or faces (config 5 has 23.7 M tetrahedra)."""
    n = nx * ny * nz
    ids = torch.arange(n, dtype=torch.int64, device=device).view(nx, ny, nz)
    keys = []
    for dx, dy, dz in KUHN_DIRECTIONS:
        u = ids[:nx - dx, :ny - dy, :nz - dz].reshape(-1)
        v = ids[dx:, dy:, dz:].reshape(-1)
        keys.append(u * n + v)
        keys.append(v * n + u)
    k = torch.sort(torch.cat(keys)).values
    return torch.stack([k // n, k % n])


def grid_for_nodes(target_nodes: float) -> int:
    """Edge length of the cubic grid whose node count is closest to `target_nodes`."""
    g = max(2, round(target_nodes ** (1.0 / 3.0)))
    return min((g - 1, g, g + 1), key=lambda q: abs(q ** 3 - target_nodes)) if g > 2 else g


# ---------------------------------------------------------------------------------------------------------------
# BASELINE.json config 3: DeformingPlate-shaped samples: per sample a thin plate (18 x 18 x 3 nodes)
# and an obstacle block (7 x 7 x 6 nodes) hovering over it, both Kuhn-split tetrahedral grids; node types NORMAL (0,
# plate), HANDLE (3, the plate's x = 0 face) and OBSTACLE (1).  The obstacle sits `gap` above the plate so that its
# bottom nodes come within the world-edge radius (0.03, preprocessing.py:92-96) of plate nodes: ~1-3 % extra edges.
PLATE_DIMS, OBSTACLE_DIMS = (18, 18, 3), (7, 7, 6)
NODE_NORMAL, NODE_OBSTACLE, NODE_HANDLE = 0, 1, 3


def _grid_points(dims, spacing, origin, gen, jitter=0.2):
    nx, ny, nz = dims
    ix, iy, iz = torch.meshgrid(torch.arange(nx), torch.arange(ny), torch.arange(nz), indexing="ij")
    p = torch.stack([ix.flatten(), iy.flatten(), iz.flatten()], dim=1).double()
    sp = torch.tensor(spacing, dtype=torch.float64)
    p = p * sp + torch.tensor(origin, dtype=torch.float64)
    p = p + (torch.rand(p.shape, generator=gen, dtype=torch.float64) * 2 - 1) * jitter * sp
    return p


def plate_sample(seed: int = 0, gap: float = 0.018):
    """One config-3 sample on the CPU: (pos [n,3] fp32 = mesh = world positions, cells [T,4] int64, node_type [n]
    int64).  n = 972 + 294 = 1,266 nodes, T = 6 (17*17*2 + 6*6*5) = 4,548 tetrahedra."""
    gen = torch.Generator().manual_seed(1234 + seed)
    psp, osp = (0.025, 0.025, 0.010), (0.020, 0.020, 0.020)
    plate = _grid_points(PLATE_DIMS, psp, (0.0, 0.0, 0.0), gen)
    top = psp[2] * (PLATE_DIMS[2] - 1)
    off = torch.rand(2, generator=gen, dtype=torch.float64) * 0.2          # where over the plate the obstacle hangs
    obst = _grid_points(OBSTACLE_DIMS, osp, (0.05 + float(off[0]), 0.05 + float(off[1]), top + gap), gen, jitter=0.1)
    n_p = plate.shape[0]
    cells = torch.cat([kuhn_grid_cells(*PLATE_DIMS), kuhn_grid_cells(*OBSTACLE_DIMS) + n_p], dim=0)
    node_type = torch.cat([torch.full((n_p,), NODE_NORMAL), torch.full((obst.shape[0],), NODE_OBSTACLE)]).long()
    ids = torch.arange(n_p).view(*PLATE_DIMS)
    node_type[ids[0].reshape(-1)] = NODE_HANDLE
    return torch.cat([plate, obst]).float(), cells, node_type


def config3_batch(copies: int = 16, device="cuda", radius: float = 0.03):
    """BASELINE.json config 3 (plate.json graph: mesh + world edges, 8-dim raw edge features) as one block-diagonal
    batch built with the device set-up path the dataset uses (cells_to_faces -> face_to_edge -> add_world_edges per
    sample -> edge_features + world_pos_features; graphphysics/dataset/preprocessing.py:401-427).
    Returns dict(edge_index [2,E], num_nodes, edge_attr [E,8], node_type [N], pos [N,3], mesh_edges, world_edges)."""
    from . import graph as G
    eis, poss, types = [], [], []
    n_off = mesh_e = 0
    for s in range(copies):
        pos, cells, nt = plate_sample(seed=s, gap=0.012 + 0.014 * ((s * 7) % copies) / max(copies - 1, 1))
        pos, cells, nt = pos.to(device), cells.to(device), nt.to(device)
        n = pos.shape[0]
        ei_mesh = G.face_to_edge(G.cells_to_faces(cells), n)
        ei = G.add_world_edges(ei_mesh, pos, nt, n, radius=radius, obstacle=NODE_OBSTACLE, normal=NODE_NORMAL)
        mesh_e += ei_mesh.shape[1]
        eis.append(ei + n_off)
        poss.append(pos)
        types.append(nt)
        n_off += n
    ei = torch.cat(eis, dim=1).contiguous()
    pos = torch.cat(poss)
    ea = G.world_pos_features(G.edge_features(pos, ei), pos, ei)
    return {"edge_index": ei, "num_nodes": n_off, "edge_attr": ea, "node_type": torch.cat(types), "pos": pos,
            "mesh_edges": mesh_e, "world_edges": ei.shape[1] - mesh_e}


def config4_graph(grid: int = 47, device="cuda", num_hops: int = 2, masking_ratio: float = 0.4, seed: int = 0):
    """BASELINE.json config 4: grid^3 Kuhn tetrahedral mesh (47^3 = 103,823 nodes), 2-hop adjacency
    (compute_k_hop_edge_index, graphphysics/utils/torch_graph.py:16-53) and the MeshMask sub-graph of the kept nodes
    (get_masked_indexes + build_masked_graph, torch_graph.py:285-302, utils/meshmask.py:9-70; masking_ratio 0.4 as
    tests/graphphysics/dataset/test_h5dataset.py:34, torch.manual_seed(seed) randperm).
    Returns dict(full=(edge_index, N), masked=(edge_index, N_kept), selected)."""
    from . import graph as G
    from .dataset import masked_indexes
    n = grid ** 3
    ei1 = kuhn_grid_edges(grid, grid, grid, device=device)
    eik = G.compute_k_hop_edge_index(ei1, num_hops, n)
    gen = torch.Generator().manual_seed(seed)
    sel = masked_indexes(n, masking_ratio, generator=gen).to(device)
    ei_m, _, _ = G.filter_edges(eik, sel)
    return {"full": (eik, n), "masked": (ei_m.contiguous(), int(sel.numel())), "selected": sel, "one_hop_edges": ei1.shape[1]}
