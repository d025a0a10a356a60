#!/usr/bin/env python
"""Fixture for the trajectory -> frame -> graph step of the dataset readers: runs the REFERENCE's
own `get_traj_as_meshes` / `get_frame_as_mesh` / `get_frame_as_graph` (graphphysics/utils/hierarchical.py:38-154) and
`meshdata_to_graph` (utils/torch_graph.py:114-193) from /root/reference on two synthetic trajectories laid out as an H5
group holds them (flat per-feature arrays + the meta.json of tests/mock_h5) and writes tests/golden/ref_dataset.npz.

    python tests/golden/make_golden_dataset.py

h5py and torch_geometric are not installed: the stand-ins of make_golden.py are registered, plus an empty `h5py` module
(the functions above only index the handle: `handle[traj][key][()]`, which a dict of numpy arrays answers the same way).
Build container only (reads /root/reference); the tests use the committed output.
"""
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, _install_stubs  # noqa: E402

T_2D, T_3D = 6, 4


def trajectories():
    """{name: (group, meta)}: `group` = feature name -> array as stored (flattened), `meta` = the dataset's meta.json."""
    rng = np.random.default_rng(42)
    z = np.load(os.path.join(HERE, "cylinder_mesh.npz"))
    pts, cells = z["points"][:, :2].astype(np.float32), z["cells"].astype(np.int32)
    n = pts.shape[0]
    # (the reference indexes `mesh_pos` and `cells` by frame whenever they have a frame axis, hierarchical.py:118-121:
    # statics are stored once per frame, as the MeshGraphNets datasets do; only node_type is read at frame 0)
    g2 = {"cells": np.tile(cells.reshape(-1), T_2D),
          "mesh_pos": np.tile(pts.reshape(-1), T_2D).astype(np.float64),        # stored wider: astype path
          "node_type": rng.integers(0, 7, size=(n,)).astype(np.int64),
          "velocity": rng.standard_normal((T_2D, n, 2)).astype(np.float32).reshape(T_2D, -1),
          "pressure": rng.standard_normal((T_2D * n,)).astype(np.float32)}
    m2 = {"dt": 0.01, "trajectory_length": T_2D, "features": {
        "cells": {"type": "static", "shape": [T_2D, -1, 3], "dtype": "int32"},
        "mesh_pos": {"type": "static", "shape": [T_2D, -1, 2], "dtype": "float32"},
        "node_type": {"type": "static", "shape": [1, -1, 1], "dtype": "int32"},
        "velocity": {"type": "dynamic", "shape": [T_2D, -1, 2], "dtype": "float32"},
        "pressure": {"type": "dynamic", "shape": [T_2D, -1, 1], "dtype": "float32"}}}
    # 3-D: tetrahedra, positions and cells stored per frame, explicit target features, no dt
    za = np.load(os.path.join(HERE, "aneurysm_mesh.npz"))
    tets = za["cells"][:150].astype(np.int64)
    used, inv = np.unique(tets.reshape(-1), return_inverse=True)
    tets = inv.reshape(-1, 4).astype(np.int32)
    p3 = za["points"][used].astype(np.float32)
    n3 = p3.shape[0]
    g3 = {"cells": np.broadcast_to(tets, (T_3D,) + tets.shape).copy(),
          "mesh_pos": (p3[None] + 0.01 * rng.standard_normal((T_3D, n3, 3))).astype(np.float32),
          "node_type": rng.integers(0, 7, size=(1, n3, 1)).astype(np.int32),
          "Vitesse": rng.standard_normal((T_3D, n3, 3)).astype(np.float32),
          "wall_mask": rng.integers(0, 2, size=(T_3D, n3, 1)).astype(np.float32)}
    m3 = {"trajectory_length": T_3D, "target_features": ["Vitesse"], "features": {
        "cells": {"type": "static", "shape": [T_3D, -1, 4], "dtype": "int32"},
        "mesh_pos": {"type": "static", "shape": [T_3D, -1, 3], "dtype": "float32"},
        "node_type": {"type": "static", "shape": [1, -1, 1], "dtype": "int32"},
        "Vitesse": {"type": "dynamic", "shape": [T_3D, -1, 3], "dtype": "float32"},
        "wall_mask": {"type": "static", "shape": [T_3D, -1, 1], "dtype": "float32"}}}
    return {"cyl2d": (g2, m2), "tet3d": (g3, m3)}


CASES = {"cyl2d": [(0, 1), (3, 4), (4, None), (T_2D - 2, T_2D - 1)], "tet3d": [(0, 1), (2, 3), (1, None)]}


def main():
    _install_stubs()
    sys.modules.setdefault("h5py", types.ModuleType("h5py"))
    sys.modules["h5py"].File = object
    sys.path.insert(0, REF)
    from graphphysics.utils.hierarchical import get_frame_as_graph, get_traj_as_meshes
    from graphphysics.utils.torch_graph import meshdata_to_graph

    out = {}
    for name, (group, meta) in trajectories().items():
        out[f"{name}.meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
        for k, v in group.items():
            out[f"{name}.raw.{k}"] = v
        traj = get_traj_as_meshes({"t0": group}, "t0", meta)
        for frame, target in CASES[name]:
            g = get_frame_as_graph(traj, frame, meta, frame_target=target)
            tag = f"{name}.f{frame}_{'n' if target is None else target}"
            out[f"{tag}.x"] = g.x.numpy()
            out[f"{tag}.pos"] = g.pos.numpy()
            out[f"{tag}.face"] = g.face.numpy()
            if g.tetra is not None:
                out[f"{tag}.tetra"] = g.tetra.numpy()
            if g.y is not None:
                out[f"{tag}.y"] = g.y.numpy()
    # the all-1-D branch of meshdata_to_graph (torch_graph.py:146-151, 163-165): features stacked as columns
    rng = np.random.default_rng(7)
    pts = rng.standard_normal((11, 2)).astype(np.float32)
    cells = rng.integers(0, 11, size=(5, 3)).astype(np.int32)
    pd = {"a": rng.standard_normal(11).astype(np.float32), "b": rng.standard_normal(11).astype(np.float32)}
    tg = {"a": rng.standard_normal(11).astype(np.float32)}
    g = meshdata_to_graph(points=pts, cells=cells, point_data=pd, time=0.25, target=tg)
    out.update({"flat.points": pts, "flat.cells": cells, "flat.a": pd["a"], "flat.b": pd["b"], "flat.ta": tg["a"],
                "flat.x": g.x.numpy(), "flat.y": g.y.numpy(), "flat.face": g.face.numpy()})
    g = meshdata_to_graph(points=pts, cells=cells, point_data=None)
    out["flat.x_none"] = g.x.numpy()
    np.savez_compressed(os.path.join(HERE, "ref_dataset.npz"), **out)
    print("wrote ref_dataset.npz:", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "ref_dataset.npz")), "bytes")


if __name__ == "__main__":
    main()
