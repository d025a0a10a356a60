"""CPU: `.vtu` writer / reader and the `.pvd` trajectory index (the role of meshio in
graphphysics/utils/meshio_mesh.py:10-67 and training/lightning_module.py:124-161).  The writer's files are read back by
the package's reader AND by oracle/vtu.py, an independent decoder that was validated on the reference's meshio-written
fixtures (tests/mock_vtu/*.vtu -> tests/golden/cylinder_mesh.npz)."""
import os

import numpy as np
import pytest
import torch

from graph_physics_b200 import vtu as V
from graph_physics_b200.models import Graph
from oracle import vtu as oracle_vtu


def _graph(golden_dir, tets=False):
    g = torch.Generator().manual_seed(4)
    if tets:
        z = np.load(os.path.join(golden_dir, "aneurysm_mesh.npz"))
        cells = torch.as_tensor(z["cells"][:500].astype(np.int64))
        pos = torch.as_tensor(z["points"])
        n = pos.shape[0]
        c = cells.t()
        face = torch.cat([c[0:3], c[1:4], torch.stack([c[2], c[3], c[0]]), torch.stack([c[3], c[0], c[1]])], dim=1)
        return Graph(x=torch.randn(n, 4, generator=g), y=torch.randn(n, 3, generator=g), pos=pos, face=face, tetra=c)
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    pos = torch.as_tensor(z["points"][:, :2])
    n = pos.shape[0]
    return Graph(x=torch.randn(n, 5, generator=g), y=torch.randn(n, 2, generator=g), pos=pos,
                 face=torch.as_tensor(z["face"].astype(np.int64)), tetra=None)


@pytest.mark.parametrize("tets", [False, True])
def test_write_then_read_round_trip_and_oracle_reader(golden_dir, tmp_path, tets):
    g = _graph(golden_dir, tets)
    path = os.path.join(tmp_path, "out", "frame.vtu")
    V.write_vtu(path, g, add_all_data=True)
    for reader in (V.read_vtu, oracle_vtu.read_vtu):
        points, cells, pd = reader(path)
        assert points.shape == (g.pos.shape[0], 3)
        assert np.array_equal(points[:, :g.pos.shape[1]], g.pos.numpy())
        assert not points[:, g.pos.shape[1]:].any()                       # 2-D positions padded with zeros
        want_cells = (g.tetra if tets else g.face).numpy().T
        assert np.array_equal(cells, want_cells)
        assert sorted(pd) == sorted([f"x{i}" for i in range(g.x.shape[1])] + [f"y{i}" for i in range(g.y.shape[1])])
        for i in range(g.x.shape[1]):
            assert np.array_equal(pd[f"x{i}"], g.x[:, i].numpy())
        for i in range(g.y.shape[1]):
            assert np.array_equal(pd[f"y{i}"], g.y[:, i].numpy())
    # without add_all_data: geometry only (meshio_mesh.py:55-66)
    V.write_vtu(path, g)
    assert V.read_vtu(path)[2] == {}


def test_block_boundaries_and_ascii_arrays(tmp_path):
    """Arrays of exactly one / two compression blocks, one byte over, and empty point data names; an ascii file."""
    for n in (8192, 8193, 16384, 1):              # float32: 8192 values = one 32 KiB block exactly
        pts = np.zeros((n, 3), dtype=np.float32)
        pts[:, 0] = np.arange(n)
        cells = np.stack([np.arange(n), (np.arange(n) + 1) % n, (np.arange(n) + 2) % n], 1).astype(np.int64)
        path = os.path.join(tmp_path, f"b{n}.vtu")
        V.write_vtu_arrays(path, pts, cells, {"s": np.arange(n, dtype=np.float32), "v": np.ones((n, 2), np.float64),
                                              "k": np.arange(n, dtype=np.int32)})
        for reader in (V.read_vtu, oracle_vtu.read_vtu):
            p, c, pd = reader(path)
            assert np.array_equal(p, pts) and np.array_equal(c, cells)
            assert np.array_equal(pd["s"], np.arange(n)) and pd["v"].shape == (n, 2) and pd["k"].dtype == np.int32
    ascii_vtu = """<?xml version="1.0"?>
<VTKFile type="UnstructuredGrid" version="0.1" byte_order="LittleEndian">
<UnstructuredGrid>
<Piece NumberOfPoints="4" NumberOfCells="2">
<Points><DataArray type="Float32" Name="Points" NumberOfComponents="3" format="ascii">
0 0 0  1 0 0  0 1 0  1 1 0
</DataArray></Points>
<Cells><DataArray type="Int64" Name="connectivity" format="ascii">0 1 2 1 3 2</DataArray>
<DataArray type="Int64" Name="offsets" format="ascii">3 6</DataArray>
<DataArray type="UInt8" Name="types" format="ascii">5 5</DataArray></Cells>
<PointData><DataArray type="Float64" Name="p" format="ascii">1.5 2.5 3.5 4.5</DataArray></PointData>
</Piece></UnstructuredGrid></VTKFile>"""
    path = os.path.join(tmp_path, "a.vtu")
    open(path, "w").write(ascii_vtu)
    p, c, pd = V.read_vtu(path)
    assert p.shape == (4, 3) and np.array_equal(c, [[0, 1, 2], [1, 3, 2]]) and np.array_equal(pd["p"], [1.5, 2.5, 3.5, 4.5])


def test_errors_and_trajectory_index(golden_dir, tmp_path):
    g = _graph(golden_dir)
    with pytest.raises(ValueError):
        V.write_vtu(os.path.join(tmp_path, "x.vtu"), Graph(x=g.x, face=g.face, pos=None))
    with pytest.raises(ValueError):
        V.write_vtu(os.path.join(tmp_path, "x.vtu"), Graph(x=g.x, face=None, pos=g.pos, tetra=None))
    with pytest.raises(ValueError):
        V.write_vtu(os.path.join(tmp_path, "x.vtu"), Graph(x=g.x, face=g.face, pos=torch.zeros(g.pos.shape[0], 4)))
    with pytest.raises(ValueError):
        V.write_vtu_arrays(os.path.join(tmp_path, "x.vtu"), np.zeros((3, 3), np.float32), np.zeros((1, 5), np.int64))
    frames = []
    for t in range(3):
        f = g.clone()
        f.x = g.x + t
        frames.append(f)
    pvd = V.write_trajectory(os.path.join(tmp_path, "pred"), "graph_0", frames, timestep=0.01)
    text = open(pvd).read()
    assert text.count("<DataSet") == 3 and 'timestep="0.01"' in text and 'timestep="0.03"' in text
    for t in range(3):
        _, _, pd = V.read_vtu(os.path.join(tmp_path, "pred", f"graph_0_{t:05d}.vtu"))
        assert np.array_equal(pd["x0"], (g.x[:, 0] + t).numpy())
