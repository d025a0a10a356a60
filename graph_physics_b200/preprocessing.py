"""Per-sample preprocessing pipeline of the reference (`graphphysics/dataset/preprocessing.py`): the
graph-object transforms that run on the device between the dataset and the model, composed in the reference's order by
`build_preprocessing` (preprocessing.py:364-446):

    [extra node features...] ->
      world positions given:  add_obstacles_next_pos, FaceToEdge, add_world_edges, Cartesian + Distance,
                              add_world_pos_features
      otherwise:              FaceToEdge, (Cartesian + Distance if add_edges_features)
    with add_noise inserted as the SECOND transform when noise parameters are given -> [extra edge features...]

Index work (FaceToEdge, world-edge radius search, edge features) is the device code of graph.py; the feature transforms
(`add_noise`, `add_obstacles_next_pos`) are written here without boolean-index assignments or in-place writes into
`graph.x`: the reference's `noise[mask] = 0` / `x[:, s:e] = ...` become a multiply by the node-type mask and a fresh
feature matrix, so (a) nothing is read back on the host and (b) a sample that is a VIEW of a device-resident trajectory
(dataset.ResidentTrajectory) is never modified.  With the same torch seed `add_noise` draws the same numbers as the
reference (one `randn_like` per feature range, in order).

Transforms take and return a duck-typed graph (`models.Graph`): `.x`, `.y`, `.pos`, `.face`, `.edge_index`, `.edge_attr`.
A graph that already carries `edge_index` (and `edge_attr`) for its mesh -- samples of `dataset.TrajectoryDataset` do,
built once per trajectory -- keeps them: FaceToEdge / Cartesian + Distance are then no-ops.
"""
from __future__ import annotations

import math
from functools import partial
from typing import Callable, List, Optional, Sequence, Union

import torch

from .utils.nodetype import NodeType

Transform = Callable[[object], object]


def _as_list(v) -> list:
    return list(v) if isinstance(v, (list, tuple)) else [v]


def add_noise(graph, noise_index_start: Union[int, Sequence[int]], noise_index_end: Union[int, Sequence[int]],
              noise_scale: Union[float, Sequence[float]], node_type_index: int, t: Optional[float] = None):
    """preprocessing.py:177-236: Gaussian noise of the given scale(s) on the feature ranges of NORMAL nodes; with `t`
    the scale follows the curriculum 10 * scale * (1 + cos(pi t))."""
    starts, ends = _as_list(noise_index_start), _as_list(noise_index_end)
    scales = [noise_scale] * len(starts) if isinstance(noise_scale, float) else _as_list(noise_scale)
    if len(starts) != len(ends):
        raise ValueError("noise_index_start and noise_index_end must have the same length.")
    if len(scales) != len(starts):
        raise ValueError("noise_scale must have the same length as noise_index_start and noise_index_end.")
    x = graph.x
    keep = (x[:, node_type_index] == int(NodeType.NORMAL)).to(x.dtype).unsqueeze(-1)      # 1 on NORMAL nodes
    out = x.clone()
    for start, end, scale in zip(starts, ends, scales):
        scale_ = 10 * scale * (1 + math.cos(t * math.pi)) if t is not None else scale
        feature = out[:, start:end]
        noise = torch.randn_like(feature) * scale_
        out[:, start:end] = feature + noise * keep
    graph.x = out
    return graph


def add_obstacles_next_pos(graph, world_pos_index_start: int, world_pos_index_end: int, node_type_index: int):
    """preprocessing.py:47-89: x <- [world_pos | obstacle displacement | rest], where the displacement target - current
    of OBSTACLE nodes is kept and every other node gets the obstacles' mean displacement.  `node_type_index` is the
    index AFTER the three displacement columns are inserted (the reference reads column `node_type_index - 3`)."""
    x, y = graph.x, graph.y
    world_pos = x[:, world_pos_index_start:world_pos_index_end]
    other = x[:, world_pos_index_end:]
    disp = y[:, world_pos_index_start:world_pos_index_end] - world_pos
    is_obst = (x[:, node_type_index - 3] == int(NodeType.OBSTACLE)).unsqueeze(-1)
    w = is_obst.to(disp.dtype)
    mean = (disp * w).sum(dim=0) / w.sum()                  # NaN without obstacles, as torch.mean of an empty set
    graph.x = torch.cat([world_pos, torch.where(is_obst, disp, mean.expand_as(disp)), other], dim=1)
    return graph


def face_to_edge(graph):
    """T.FaceToEdge(remove_faces=False) (preprocessing.py:410,429); kept if the sample already has its mesh edges."""
    if getattr(graph, "edge_index", None) is None:
        from . import graph as G
        graph.edge_index = G.face_to_edge(graph.face.contiguous(), graph.x.shape[0])
        graph.edge_attr = None
    return graph


def add_world_edges(graph, world_pos_index_start: int, world_pos_index_end: int, node_type_index: int,
                    radius: float = 0.03):
    """preprocessing.py:92-140 with the cKDTree search replaced by the device cell list (graph.world_pairs)."""
    from . import graph as G
    x = graph.x
    graph.edge_index = G.add_world_edges(graph.edge_index, x[:, world_pos_index_start:world_pos_index_end].contiguous(),
                                         x[:, node_type_index], x.shape[0], radius=radius)
    graph.edge_attr = None                  # the edge set changed: mesh-only features no longer line up
    return graph


def edge_features(graph):
    """T.Cartesian(norm=False) then T.Distance(norm=False) (preprocessing.py:16-23)."""
    if getattr(graph, "edge_attr", None) is None:
        from . import graph as G
        graph.edge_attr = G.edge_features(graph.pos.contiguous(), graph.edge_index)
    return graph


def add_world_pos_features(graph, world_pos_index_start: int, world_pos_index_end: int):
    """preprocessing.py:143-174."""
    from . import graph as G
    graph.edge_attr = G.world_pos_features(graph.edge_attr,
                                           graph.x[:, world_pos_index_start:world_pos_index_end].contiguous(),
                                           graph.edge_index)
    return graph


def compute_min_distance_to_type(graph, target_type: int, node_types: torch.Tensor) -> torch.Tensor:
    """preprocessing.py:239-271: for every node the Euclidean distance to the nearest node of `target_type` ([N]).
    The reference materialises an [N, M, 3] difference tensor; `cdist` without the matmul shortcut computes the same
    sqrt(sum(diff^2)) in [N, M]."""
    pos = graph.pos
    targets = pos[node_types == int(target_type)]
    return torch.cdist(pos, targets, compute_mode="donot_use_mm_for_euclid_dist").min(dim=1)[0]


class Random3DRotate:
    """preprocessing.py:274-366: one random rotation R (angles uniform in +-180 degrees from Python's `random`, the
    reference's composition of the three axis rotations) applied as `v @ R` to the positions, to every listed
    3-column feature range of x and to the first three target columns -- `y` is REPLACED by those three rotated columns,
    as in the reference.  x is not modified in place (a sample may be a view of a resident trajectory)."""

    def __init__(self, feature_indices=None) -> None:
        self.feature_indices = feature_indices or []
        for start_idx, end_idx in self.feature_indices:
            assert end_idx - start_idx == 3, (
                f"Each feature range must span exactly 3 features for xyz coordinates. Got range {start_idx}-{end_idx}")

    def _get_random_angles(self):
        import random
        return [math.radians(random.uniform(-180, 180)) for _ in range(3)]

    @staticmethod
    def _build_rotation_matrix(alpha: float, beta: float, gamma: float) -> torch.Tensor:
        ca, sa, cb, sb, cg, sg = (math.cos(alpha), math.sin(alpha), math.cos(beta), math.sin(beta), math.cos(gamma),
                                  math.sin(gamma))
        return torch.tensor([[ca * cb, ca * sb * sg + sa * cg, -ca * sb * cg + sa * sg],
                             [-sa * cb, -sa * sb * sg + ca * cg, sa * sb * cg + ca * sg],
                             [sb, -cb * sg, cb * cg]])

    def __call__(self, data):
        return self.forward(data)

    def forward(self, data):
        matrix = self._build_rotation_matrix(*self._get_random_angles())
        if getattr(data, "pos", None) is not None:
            pos = data.pos.view(-1, 1) if data.pos.dim() == 1 else data.pos
            assert pos.size(-1) == 3, "Node positions must be 3-dimensional"
            data.pos = pos @ matrix.to(pos.device, pos.dtype)
        if getattr(data, "x", None) is not None and self.feature_indices:
            x = data.x.clone()
            for start_idx, end_idx in self.feature_indices:
                x[:, start_idx:end_idx] = x[:, start_idx:end_idx] @ matrix.to(x.device, x.dtype)
            data.x = x
        if getattr(data, "y", None) is not None and getattr(data, "x", None) is not None:
            target = data.y[:, 0:3]
            data.y = target @ matrix.to(target.device, target.dtype)
        return data


class Compose:
    """T.Compose: applies the transforms in order; `.transforms` is the list (as PyG exposes it)."""

    def __init__(self, transforms: List[Transform]):
        self.transforms = list(transforms)

    def __call__(self, graph):
        for t in self.transforms:
            graph = t(graph)
        return graph


def build_preprocessing(noise_parameters: Optional[dict] = None, world_pos_parameters: Optional[dict] = None,
                        add_edges_features: bool = True, extra_node_features=None, extra_edge_features=None) -> Compose:
    """preprocessing.py:364-446, same arguments and the same order of transforms."""
    pre: List[Transform] = []
    if extra_node_features is not None:
        pre.extend(_as_list(extra_node_features))
    if world_pos_parameters is not None:
        w = world_pos_parameters
        rng = dict(world_pos_index_start=w["world_pos_index_start"], world_pos_index_end=w["world_pos_index_end"])
        pre.extend([partial(add_obstacles_next_pos, node_type_index=w["node_type_index"], **rng),
                    face_to_edge,
                    partial(add_world_edges, node_type_index=w["node_type_index"], radius=w.get("radius", 0.03), **rng),
                    edge_features,
                    partial(add_world_pos_features, **rng)])
    else:
        pre.append(face_to_edge)
        if add_edges_features:
            pre.append(edge_features)
    if noise_parameters is not None:
        n = noise_parameters
        pre.insert(1, partial(add_noise, noise_index_start=n["noise_index_start"], noise_index_end=n["noise_index_end"],
                              noise_scale=n["noise_scale"], node_type_index=n["node_type_index"]))
    if extra_edge_features is not None:
        pre.extend(_as_list(extra_edge_features))
    return Compose(pre)
