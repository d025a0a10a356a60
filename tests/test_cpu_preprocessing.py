"""CPU: the feature transforms of the per-sample preprocessing and the order `build_preprocessing` composes them in
 against tests/golden/ref_preprocessing.npz, written by the reference's own
graphphysics/dataset/preprocessing.py (tests/golden/make_golden_preprocessing.py)."""
import json
import os

import numpy as np
import pytest
import torch

from graph_physics_b200 import preprocessing as P
from graph_physics_b200.models import Graph


def _name(t):
    f = getattr(t, "func", t)
    return f.__name__


def test_add_noise_draws_the_reference_numbers(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_preprocessing.npz"))
    x = torch.from_numpy(z["x"])
    cases = {"one": dict(noise_index_start=0, noise_index_end=2, noise_scale=0.02),
             "two": dict(noise_index_start=[0, 3], noise_index_end=[2, 6], noise_scale=[0.02, 0.5]),
             "curriculum": dict(noise_index_start=[1], noise_index_end=[4], noise_scale=[0.1], t=0.3)}
    for name, kw in cases.items():
        torch.manual_seed(1234)
        keep = x.clone()
        g = P.add_noise(Graph(x=keep), node_type_index=7, **kw)
        assert np.array_equal(g.x.numpy(), z[f"noise_{name}"]), name
        assert torch.equal(keep, x)                                   # the input tensor (a resident view) is untouched
        changed = (g.x != x).any(dim=1)
        assert not changed[x[:, 7] != 0].any() and changed[x[:, 7] == 0].all()        # NORMAL nodes only
    with pytest.raises(ValueError):
        P.add_noise(Graph(x=x), [0, 1], [2], 0.1, 7)
    with pytest.raises(ValueError):
        P.add_noise(Graph(x=x), [0, 1], [2, 3], [0.1], 7)


def test_add_obstacles_next_pos(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_preprocessing.npz"))
    x, y = torch.from_numpy(z["x"]), torch.from_numpy(z["y"])
    g = P.add_obstacles_next_pos(Graph(x=x.clone(), y=y), world_pos_index_start=0, world_pos_index_end=3,
                                 node_type_index=10)
    assert g.x.shape == (x.shape[0], x.shape[1] + 3)
    np.testing.assert_allclose(g.x.numpy(), z["obstacles"], rtol=0, atol=2e-7)       # masked sum vs mean of a subset
    obst = x[:, 7] == 1
    assert np.array_equal(g.x[obst].numpy(), z["obstacles"][obst.numpy()])           # obstacle rows: exact


def test_build_preprocessing_order(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_preprocessing.npz"))
    orders = json.loads(bytes(z["orders"]).decode())
    wp = {"world_pos_index_start": 0, "world_pos_index_end": 3, "node_type_index": 10, "radius": 0.05}
    nz = {"noise_index_start": 0, "noise_index_end": 2, "noise_scale": 0.02, "node_type_index": 7}

    def extra_a(gr):
        return gr

    def extra_b(gr):
        return gr

    cfgs = {"plain": {}, "no_edge_features": {"add_edges_features": False}, "noise": {"noise_parameters": nz},
            "world": {"world_pos_parameters": wp}, "world_noise": {"world_pos_parameters": wp, "noise_parameters": nz},
            "extras": {"noise_parameters": nz, "extra_node_features": extra_a, "extra_edge_features": [extra_b, extra_a]}}
    ref_names = {"face_to_edge": ["FaceToEdge"], "edge_features": ["Cartesian", "Distance"]}
    for name, kw in cfgs.items():
        pipe = P.build_preprocessing(**kw)
        got = [n for t in pipe.transforms for n in ref_names.get(_name(t), [_name(t)])]
        assert got == orders[name], name
    pipe = P.build_preprocessing(world_pos_parameters=wp)
    assert pipe.transforms[2].keywords["radius"] == 0.05
    # a sample that already carries its mesh edges and features (dataset.TrajectoryDataset) passes through unchanged
    ei, ea = torch.zeros((2, 3), dtype=torch.int64), torch.zeros((3, 3))
    g = P.build_preprocessing()(Graph(x=torch.zeros(4, 2), edge_index=ei, edge_attr=ea, face=None, pos=None))
    assert g.edge_index is ei and g.edge_attr is ea


def test_min_distance_and_random_rotation(golden_dir):
    """preprocessing.py:239-366 (the extra features of the 3-D configurations)."""
    import random
    z = np.load(os.path.join(golden_dir, "ref_preprocessing.npz"))
    pos3, types = torch.from_numpy(z["pos3"]), torch.from_numpy(z["types3"])
    d = P.compute_min_distance_to_type(Graph(pos=pos3), 4, types)
    np.testing.assert_allclose(d.numpy(), z["min_dist_inflow"], rtol=1e-6, atol=0)
    assert float(d[types == 4].max()) == 0.0
    x, y = torch.from_numpy(z["x"]), torch.from_numpy(z["y"])
    random.seed(5)
    keep = x.clone()
    g = P.Random3DRotate(feature_indices=[(0, 3), (4, 7)])(Graph(x=keep, y=y.clone(), pos=pos3.clone()))
    assert np.array_equal(g.x.numpy(), z["rot_x"]) and np.array_equal(g.y.numpy(), z["rot_y"])
    assert np.array_equal(g.pos.numpy(), z["rot_pos"]) and torch.equal(keep, x) and g.y.shape[1] == 3
    # a rotation: lengths are kept
    assert torch.allclose(g.pos.norm(dim=1), pos3.norm(dim=1), rtol=1e-5)
    with pytest.raises(AssertionError):
        P.Random3DRotate(feature_indices=[(0, 2)])
