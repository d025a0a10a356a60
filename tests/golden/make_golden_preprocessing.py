#!/usr/bin/env python
"""Fixture for the feature transforms of the per-sample preprocessing: runs the REFERENCE's own
`add_noise`, `add_obstacles_next_pos`, `compute_min_distance_to_type`, `Random3DRotate` and `build_preprocessing`
(graphphysics/dataset/preprocessing.py:47-89, 177-446) from /root/reference on seeded inputs and writes tests/golden/ref_preprocessing.npz (+ the order of the
transforms `build_preprocessing` composes, as names).

    python tests/golden/make_golden_preprocessing.py

torch_geometric is not installed: the stand-ins of make_golden.py are registered first.  Build container only."""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, _install_stubs  # noqa: E402

N = 157


def inputs():
    g = torch.Generator().manual_seed(21)
    x = torch.randn(N, 8, generator=g)
    x[:, 7] = torch.randint(0, 7, (N,), generator=g).float()          # node type last
    y = torch.randn(N, 4, generator=g)
    return x, y


def _name(t):
    f = getattr(t, "func", t)
    return getattr(f, "__name__", type(f).__name__)


def main():
    Data = _install_stubs()
    sys.path.insert(0, REF)
    from graphphysics.dataset.preprocessing import (Random3DRotate, add_noise, add_obstacles_next_pos,
                                                    build_preprocessing, compute_min_distance_to_type)

    out = {}
    x, y = inputs()
    out["x"], out["y"] = x.numpy(), y.numpy()
    cases = {"one": dict(noise_index_start=0, noise_index_end=2, noise_scale=0.02),
             "two": dict(noise_index_start=[0, 3], noise_index_end=[2, 6], noise_scale=[0.02, 0.5]),
             "curriculum": dict(noise_index_start=[1], noise_index_end=[4], noise_scale=[0.1], t=0.3)}
    for name, kw in cases.items():
        torch.manual_seed(1234)
        g = add_noise(Data(x=x.clone()), node_type_index=7, **kw)
        out[f"noise_{name}"] = g.x.numpy()
    # x = [world_pos(3) | rest ... | node type]; after the call the node type sits 3 columns later: index 7 + 3
    g = add_obstacles_next_pos(Data(x=x.clone(), y=y.clone()), world_pos_index_start=0, world_pos_index_end=3,
                               node_type_index=10)
    out["obstacles"] = g.x.numpy()
    # extra features of the 3-D (aneurysm) configurations: min distance to a node type, random rigid rotation
    import random
    g3 = torch.Generator().manual_seed(8)
    pos3 = torch.randn(N, 3, generator=g3)
    types = torch.randint(0, 7, (N,), generator=g3)
    out["pos3"], out["types3"] = pos3.numpy(), types.numpy()
    out["min_dist_inflow"] = compute_min_distance_to_type(Data(pos=pos3), 4, types).numpy()
    random.seed(5)
    rot = Random3DRotate(feature_indices=[(0, 3), (4, 7)])
    gr = rot(Data(x=x.clone(), y=y.clone(), pos=pos3.clone()))
    out["rot_x"], out["rot_y"], out["rot_pos"] = gr.x.numpy(), gr.y.numpy(), gr.pos.numpy()
    orders = {}
    wp = {"world_pos_index_start": 0, "world_pos_index_end": 3, "node_type_index": 10, "radius": 0.05}
    nz = {"noise_index_start": 0, "noise_index_end": 2, "noise_scale": 0.02, "node_type_index": 7}

    def extra_a(gr):
        return gr

    def extra_b(gr):
        return gr

    for name, kw in {"plain": {}, "no_edge_features": {"add_edges_features": False}, "noise": {"noise_parameters": nz},
                     "world": {"world_pos_parameters": wp}, "world_noise": {"world_pos_parameters": wp, "noise_parameters": nz},
                     "extras": {"noise_parameters": nz, "extra_node_features": extra_a,
                                "extra_edge_features": [extra_b, extra_a]}}.items():
        orders[name] = [_name(t) for t in build_preprocessing(**kw).transforms]
    out["orders"] = np.frombuffer(json.dumps(orders).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "ref_preprocessing.npz"), **out)
    print(json.dumps(orders, indent=1))


if __name__ == "__main__":
    main()
