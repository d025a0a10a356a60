"""CPU: the HDF5 decoder (`h5lite`) and the XDMF time-series reader against the reference's own fixtures
(tests/mock_xdmf: contiguous little-endian float32 / int64, written by meshio's TimeSeriesWriter; tests/mock_airfoil:
chunked + deflate, BIG-endian float64 / int64).  The fixture files are the reference's and are not copied into this
repository: these tests read them from /root/reference in the build container and are skipped where it is absent
(they carry no `gpu` mark, so nothing on the GPU box touches /root/reference).  What they decode is compared with
independent routes to the same data: the committed mesh fixture (decoded from the reference's .vtu files by
oracle/vtu.py), the per-frame velocities decoded from those .vtu files, geometric consistency of points and cells."""
import hashlib
import os

import numpy as np
import pytest

from graph_physics_b200 import dataset as D, h5lite
from graph_physics_b200.xdmf import XdmfTimeSeries

REF = "/root/reference/tests"
needs_ref = pytest.mark.skipif(not os.path.isdir(REF), reason="the reference's fixture files are not on this machine")


@needs_ref
def test_h5lite_contiguous_file_matches_the_vtu_route(golden_dir):
    f = h5lite.File(os.path.join(REF, "mock_xdmf", "mock.h5"))
    assert f.keys() == sorted(f"data{i}" for i in range(14)) and "data3" in f and len(f) == 14
    mesh = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    d0, d1 = f["data0"], f["/data1"]
    assert d0.shape == (1923, 3) and d0.dtype == np.dtype("<f4") and d1.dtype == np.dtype("<i8")
    assert np.array_equal(d0[()], mesh["points"]) and np.array_equal(d1[()], mesh["cells"])
    vel = np.load(os.path.join(golden_dir, "cylinder_velocity_frames.npz"))["velocity"]
    for t in range(6):
        assert np.array_equal(f[f"data{2 + 2 * t}"][()], vel[t, :, 0])
        assert np.array_equal(f[f"data{3 + 2 * t}"][()], vel[t, :, 1])
    with pytest.raises(KeyError):
        f["nope"]
    # a group reached through its object header (symbol-table message) -- the route every nested group takes -- lists
    # the same links as the root reached through the superblock's cached addresses
    r = f._r
    root_header = r.addr(24 + 4 * r.O + r.O)
    again = f._object(root_header, "/")
    assert isinstance(again, h5lite.Group) and again.keys() == f.keys()
    assert np.array_equal(again["data2"][()], f["data2"][()])
    f.close()


@needs_ref
def test_h5lite_chunked_deflate_big_endian():
    f = h5lite.File(os.path.join(REF, "mock_airfoil", "sample_000000005.h5"))
    pts, cells = f["data0"][()], f["data1"][()]
    assert f["data0"]._layout[0] == "chunked" and f["data0"]._filters[0][0] == 1 and f["data0"].dtype == np.dtype(">f8")
    assert pts.dtype == np.float64 and pts.shape == (27125, 3) and cells.shape == (52656, 3)
    # regression pins of this decode
    want = {"data0": "ba22eb9abc27cf13", "data1": "822c8419260eaeb6", "data5": "2a6f47394fa7a548", "data7": "cf25bd4033e68471"}
    for k, h in want.items():
        assert hashlib.sha256(np.ascontiguousarray(f[k][()]).tobytes()).hexdigest()[:16] == h, k
    # independent consistency: a mis-assembled chunk or byte order would break the mesh
    assert np.isfinite(pts).all() and cells.min() == 0 and cells.max() == 27124 and np.unique(cells).size == 27125
    a, b, c = pts[cells[:, 0]], pts[cells[:, 1]], pts[cells[:, 2]]
    area = 0.5 * np.abs((b[:, 0] - a[:, 0]) * (c[:, 1] - a[:, 1]) - (b[:, 1] - a[:, 1]) * (c[:, 0] - a[:, 0]))
    assert (area > 0).all()
    longest = np.max([np.linalg.norm(a - b, axis=1), np.linalg.norm(b - c, axis=1), np.linalg.norm(c - a, axis=1)])
    assert longest < 0.25 * np.ptp(pts[:, 0])                 # a triangulation of the domain, not scrambled indices
    node_type = f["data7"][()]
    assert set(np.unique(node_type)) <= {0.0, 2.0, 4.0, 5.0, 6.0}
    f.close()


@needs_ref
def test_xdmf_time_series_and_dataset_frames(golden_dir):
    """tests/graphphysics/dataset/test_xdmfdataset.py:21-35 (mock_xdmf + meta10.json): 1,923 nodes, no edge_index yet;
    plus the frame content against the .vtu route."""
    import json
    mesh = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    vel = np.load(os.path.join(golden_dir, "cylinder_velocity_frames.npz"))["velocity"]
    meta = json.load(open(os.path.join(REF, "mock_h5", "meta10.json")))
    with XdmfTimeSeries(os.path.join(REF, "mock_xdmf", "mock.xdmf")) as reader:
        assert reader.num_steps == 6
        points, cells = reader.read_points_cells()
        assert np.array_equal(points, mesh["points"]) and np.array_equal(cells["triangle"], mesh["cells"])
        time, pd, cd = reader.read_data(4)
        assert time == 4.0 and list(pd) == ["velocity_x", "velocity_y"] and cd == {}
        g = D.xdmf_frame_as_graph(reader, 2, meta, mesh_id="mock")
        assert g.x.shape == (1923, 3) and g.y.shape == (1923, 2) and not hasattr(g, "edge_index")
        assert np.array_equal(g.x[:, :2].numpy(), vel[2]) and float(g.x[0, 2]) == 2.0      # [vx | vy | time]
        assert np.array_equal(g.y.numpy(), vel[3]) and g.target_dt == pytest.approx(0.01) and g.id == "mock"
        assert np.array_equal(g.face.numpy(), mesh["cells"].T) and g.tetra is None
        g = D.xdmf_frame_as_graph(reader, 4, meta, target_step=3, previous_step=2)            # target clamps to +1
        assert np.array_equal(g.y.numpy(), vel[5]) and g.target_dt == pytest.approx(0.01)
        assert np.array_equal(g.previous_data["velocity_x"][:, 0], vel[2, :, 0]) and g.previous_dt == pytest.approx(-0.02)
        with pytest.raises(IndexError):
            D.xdmf_frame_as_graph(reader, 5, meta)
    with XdmfTimeSeries(os.path.join(REF, "mock_airfoil", "sample_000000005.xdmf")) as reader:      # one uniform grid
        assert reader.num_steps == 1
        _, pd, _ = reader.read_data(0)
        assert list(pd) == ["Velocity_x", "Velocity_y", "Pressure", "Mach", "Distance", "Node_type"]


def test_xdmf_inline_data_and_h5lite_errors(tmp_path):
    """No reference files needed: inline (`Format="XML"`) heavy data; files that are not HDF5 / not supported."""
    xdmf = """<Xdmf Version="3.0"><Domain><Grid Name="Grid"><Geometry GeometryType="XY">
<DataItem DataType="Float" Dimensions="4 2" Format="XML" Precision="4">0 0 1 0 0 1 1 1</DataItem></Geometry>
<Topology TopologyType="Triangle" NumberOfElements="2"><DataItem DataType="Int" Dimensions="2 3" Format="XML" Precision="8">
0 1 2 1 3 2</DataItem></Topology>
<Attribute Name="p" AttributeType="Scalar" Center="Node"><DataItem DataType="Float" Dimensions="4" Format="XML" Precision="8">
1.5 2.5 3.5 4.5</DataItem></Attribute>
<Attribute Name="c" AttributeType="Scalar" Center="Cell"><DataItem DataType="Int" Dimensions="2" Format="XML" Precision="4">
7 8</DataItem></Attribute></Grid></Domain></Xdmf>"""
    path = os.path.join(tmp_path, "one.xdmf")
    open(path, "w").write(xdmf)
    with XdmfTimeSeries(path) as r:
        pts, cells = r.read_points_cells()
        assert pts.shape == (4, 2) and np.array_equal(cells["triangle"], [[0, 1, 2], [1, 3, 2]])
        t, pd, cd = r.read_data(0)
        assert t == 0.0 and np.array_equal(pd["p"], [1.5, 2.5, 3.5, 4.5]) and np.array_equal(cd["c"], [7, 8])
        with pytest.raises(IndexError):
            r.read_data(1)
    bad = os.path.join(tmp_path, "bad.h5")
    open(bad, "wb").write(b"not an hdf5 file at all")
    with pytest.raises(ValueError):
        h5lite.File(bad)
    newer = os.path.join(tmp_path, "v3.h5")
    open(newer, "wb").write(b"\\x89HDF\\r\\n\\x1a\\n".decode("unicode_escape").encode("latin1") + bytes([3]) + bytes(64))
    with pytest.raises(h5lite.H5Unsupported):
        h5lite.File(newer)
    with pytest.raises(ValueError):
        h5lite.File(bad, "w")


def test_xdmf_writer_round_trip(golden_dir, tmp_path):
    """Writer (binary heavy data) -> reader: the series `_save_trajectory_to_xdmf` would write for a predicted
    trajectory comes back bit-identical, times included."""
    import torch
    from graph_physics_b200.models import Graph
    from graph_physics_b200.xdmf import XdmfTimeSeriesWriter, write_trajectory
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    vel = np.load(os.path.join(golden_dir, "cylinder_velocity_frames.npz"))["velocity"]
    pos = torch.as_tensor(z["points"][:, :2])
    face = torch.as_tensor(z["face"].astype(np.int64))
    traj = [Graph(x=torch.as_tensor(vel[t]), y=torch.as_tensor(vel[t + 1]), pos=pos, face=face, tetra=None)
            for t in range(5)]
    path = write_trajectory(os.path.join(tmp_path, "pred"), "graph_epoch_0", traj, timestep=0.01)
    assert path.endswith("graph_epoch_0.xdmf") and os.path.exists(path[:-5] + ".bin")
    with XdmfTimeSeries(path) as r:
        assert r.num_steps == 5
        points, cells = r.read_points_cells()
        assert np.array_equal(points[:, :2], z["points"][:, :2]) and not points[:, 2].any()
        assert np.array_equal(cells["triangle"], z["cells"])
        for t in range(5):
            time, pd, cd = r.read_data(t)
            assert time == pytest.approx(0.01 * (t + 1)) and cd == {} and list(pd) == ["x0", "x1", "y0", "y1"]
            assert np.array_equal(pd["x1"], vel[t, :, 1]) and np.array_equal(pd["y0"], vel[t + 1, :, 0])
    with XdmfTimeSeries(write_trajectory(os.path.join(tmp_path, "pred"), "prev", traj[:2], 0.5, True)) as r:
        assert r.read_data(0)[0] == 1.0 and r.read_data(1)[0] == 1.5
    with pytest.raises(ValueError):
        XdmfTimeSeriesWriter(os.path.join(tmp_path, "x.h5"))
    with XdmfTimeSeriesWriter(os.path.join(tmp_path, "v.xdmf")) as w:
        with pytest.raises(ValueError):
            w.write_data(0.0, point_data={})
        w.write_points_cells(np.zeros((3, 3), np.float32), [("triangle", np.array([[0, 1, 2]]))])
        with pytest.raises(ValueError):
            w.write_data(0.0, point_data={"bad": np.zeros(7)})
        w.write_data(0.0, point_data={"vec": np.ones((3, 3))}, cell_data={"c": np.array([5], dtype=np.int32)})
    with XdmfTimeSeries(os.path.join(tmp_path, "v.xdmf")) as r:
        _, pd, cd = r.read_data(0)
        assert pd["vec"].shape == (3, 3) and np.array_equal(cd["c"], [5])
