"""CPU: trajectory -> frame -> graph against tests/golden/ref_dataset.npz, which
tests/golden/make_golden_dataset.py wrote by running the reference's `get_traj_as_meshes` / `get_frame_as_graph`
(graphphysics/utils/hierarchical.py:38-154) and `meshdata_to_graph` (utils/torch_graph.py:114-193).  Bit-exact: these
are casts, reshapes, slices and concatenations."""
import json
import os

import numpy as np
import pytest
import torch

from graph_physics_b200 import dataset as D

CASES = {"cyl2d": [(0, 1), (3, 4), (4, None), (4, 5)], "tet3d": [(0, 1), (2, 3), (1, None)]}


def _load(golden_dir, name):
    z = np.load(os.path.join(golden_dir, "ref_dataset.npz"))
    meta = json.loads(bytes(z[f"{name}.meta"]).decode())
    group = {k[len(name) + 5:]: z[k] for k in z.files if k.startswith(f"{name}.raw.")}
    return z, meta, group


@pytest.mark.parametrize("name", ["cyl2d", "tet3d"])
def test_frames_match_the_reference(golden_dir, name):
    z, meta, group = _load(golden_dir, name)
    traj = D.traj_as_meshes(group, meta)
    assert list(traj.keys()) == list(meta["features"].keys())
    for k, f in meta["features"].items():
        assert traj[k].dtype == np.dtype(f["dtype"]) and traj[k].ndim == 3
    for frame, target in CASES[name]:
        g = D.frame_as_graph(traj, frame, meta, frame_target=target)
        tag = f"{name}.f{frame}_{'n' if target is None else target}"
        assert g.x.dtype == torch.float32 and np.array_equal(g.x.numpy(), z[f"{tag}.x"])
        assert np.array_equal(g.pos.numpy(), z[f"{tag}.pos"])
        assert np.array_equal(g.face.numpy(), z[f"{tag}.face"])
        if name == "tet3d":
            assert np.array_equal(g.tetra.numpy(), z[f"{tag}.tetra"])
            assert g.face.shape[1] == 4 * g.tetra.shape[1]
        else:
            assert g.tetra is None
        if target is None:
            assert g.y is None
        else:
            assert np.array_equal(g.y.numpy(), z[f"{tag}.y"])


@pytest.mark.parametrize("name", ["cyl2d", "tet3d"])
def test_whole_trajectory_tensors_equal_the_per_frame_graphs(golden_dir, name):
    """What stays resident on the device: x[t], y[t + 1] must be frame t's graph exactly."""
    _, meta, group = _load(golden_dir, name)
    traj = D.traj_as_meshes(group, meta)
    x, y, pos, cells = D.trajectory_tensors(traj, meta)
    T = meta["trajectory_length"]
    assert x.shape[0] == T and y.shape[0] == T
    for t in range(T - 1):
        g = D.frame_as_graph(traj, t, meta, frame_target=t + 1)
        assert torch.equal(x[t], g.x) and torch.equal(y[t + 1], g.y)
        assert torch.equal(pos[t if pos.shape[0] > 1 else 0], g.pos)
        face, _ = D.faces_of_cells(cells[t if cells.shape[0] > 1 else 0])
        assert torch.equal(face, g.face)


def test_flat_point_data_and_missing_point_data(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_dataset.npz"))
    g = D.meshdata_to_graph(points=z["flat.points"], cells=z["flat.cells"],
                            point_data={"a": z["flat.a"], "b": z["flat.b"]}, time=0.25, target={"a": z["flat.ta"]})
    assert np.array_equal(g.x.numpy(), z["flat.x"]) and np.array_equal(g.y.numpy(), z["flat.y"])
    assert np.array_equal(g.face.numpy(), z["flat.face"])
    g = D.meshdata_to_graph(points=z["flat.points"], cells=z["flat.cells"], point_data=None)
    assert np.array_equal(g.x.numpy(), z["flat.x_none"]) and g.y is None
    only_x = D.meshdata_to_graph(points=z["flat.points"], cells=z["flat.cells"],
                                 point_data={"a": z["flat.a"]}, return_only_node_features=True)
    assert torch.is_tensor(only_x) and only_x.shape == (11, 2)
    with pytest.raises(ValueError):
        D.faces_of_cells(torch.zeros((4, 5), dtype=torch.int64))


def test_npz_container_and_index_arithmetic(golden_dir, tmp_path):
    _, meta, group = _load(golden_dir, "cyl2d")
    path = os.path.join(tmp_path, "train.npz")
    D.save_npz_container(path, {"traj_b": group, "traj_a": {k: v.copy() for k, v in group.items()}})
    c = D.open_container(path)
    assert c.keys() == ["traj_b", "traj_a"]                      # file order, not sorted
    got = D.traj_as_meshes(c["traj_a"], meta)
    want = D.traj_as_meshes(group, meta)
    assert all(np.array_equal(got[k], want[k]) for k in want)
    c.close()
    with pytest.raises(OSError):
        D.open_container(os.path.join(tmp_path, "train.h5"))     # no such file: the .h5 route (h5py or h5lite) says so
    with pytest.raises(ValueError):
        D.open_container(os.path.join(tmp_path, "train.xyz"))
    # dataset.py:66-85 of the reference: T - 1 samples per trajectory, shifted by one with previous data
    meta_path = os.path.join(tmp_path, "meta.json")
    json.dump(meta, open(meta_path, "w"))
    ds = D.TrajectoryDataset(path, meta_path, device="cpu")
    T = meta["trajectory_length"]
    assert ds.size_dataset == 2 and len(ds) == 2 * (T - 1)
    assert [ds.get_traj_frame(i) for i in (0, T - 2, T - 1, 2 * T - 3)] == [(0, 0), (0, T - 2), (1, 0), (1, T - 2)]
    ds.use_previous_data = True
    assert ds.get_traj_frame(0) == (0, 1) and ds.get_traj_frame(T - 1) == (1, 1)
    with pytest.raises(IndexError):
        ds[len(ds)]
    with pytest.raises(RuntimeError):
        ds[0]                 # the resident form needs the device library: no CPU path
    ds.close()


def test_masked_indexes_match_the_reference_formula():
    """utils/torch_graph.py:285-302: int((1 - ratio) n) entries of randperm(n) from the global torch generator."""
    for n, ratio in ((1923, 0.15), (10, 0.5), (7, 0.0), (5, 1.0)):
        torch.manual_seed(3)
        want = torch.randperm(n)[: int((1 - ratio) * n)]
        torch.manual_seed(3)
        got = D.masked_indexes(n, ratio)
        assert torch.equal(got, want) and got.unique().numel() == got.numel()
