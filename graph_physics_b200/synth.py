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
