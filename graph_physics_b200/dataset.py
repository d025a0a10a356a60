"""Trajectory datasets: the reference's on-disk layout -- one group per trajectory, one flat array
per feature, a `meta.json` that gives each feature's dtype, `[frames, -1, width]` shape and static / dynamic kind
(graphphysics/utils/hierarchical.py:11-154, dataset/h5_dataset.py:15-119, dataset/dataset.py:66-85) -- read into graphs
with the reference's frame semantics, and kept RESIDENT on the device.

What the reference does per sample (`H5Dataset.__getitem__`, h5_dataset.py:66-114): read EVERY feature of the whole
trajectory from the file, cast and reshape it, cut one frame out on the host, build the node-feature matrix with numpy,
copy it to the device, then run FaceToEdge / edge features / K-hop there.  For a 600-frame trajectory that is 600x the
bytes a sample needs, per sample, on the host.

Here a trajectory is decoded once: its per-frame node features and targets are assembled into `[T, N, F]` device tensors
(a 600 x 1,923-node cylinder trajectory is 23 MB of the 180 GB), its topology goes through the device graph set-up once
(`graph.face_to_edge` -> `edge_index`, `graph.edge_features`, the receiver-sorted CSR via `processor.csr_for`), and frame
`t` of the dataset is a pair of views `x[t]`, `y[t]` plus the shared `edge_index` tensor object -- no host work, no
host -> device copy and no sort per sample; `batch.collate(samples, csrs=...)` then assembles a batch's CSR from the
cached ones.  A least-recently-used budget (bytes) bounds what stays resident.

Layers, bottom up:
  * `traj_as_meshes`, `frame_as_mesh`, `meshdata_to_graph`, `frame_as_graph`: the reference's host functions, same
    results (pinned by tests/golden/ref_dataset.npz, which the reference's own code wrote), numpy / CPU torch only;
  * `open_container(path)`: trajectory groups from a `.npz` (keys `"<trajectory>/<feature>"`) or an `.h5` file --
    through h5py when it is importable, else through `h5lite`, this package's read-only decoder of the subset of the
    format h5py writes by default (anything else raises `H5Unsupported`); `xdmf_frame_as_graph` = the XDMF dataset's
    frame semantics over `xdmf.XdmfTimeSeries`;
  * `ResidentTrajectory`, `TrajectoryDataset`: the device-resident form and the `__len__` / `get_traj_frame` /
    `__getitem__` surface of the reference's `BaseDataset` (device kernels: needs the B200 library).
"""
from __future__ import annotations

import json
from collections import OrderedDict
from typing import Any, Dict, Mapping, Optional, Tuple

import numpy as np
import torch

from .models.processors import Graph

_TOPOLOGY_KEYS = ("mesh_pos", "cells", "node_type")


# ---------------------------------------------------------------------------------------------------- host functions
def load_meta(meta_path: str) -> Dict[str, Any]:
    with open(meta_path, "r") as fp:
        return json.load(fp)


def traj_as_meshes(group: Mapping[str, Any], meta: Dict[str, Any]) -> Dict[str, np.ndarray]:
    """`get_traj_as_meshes` (hierarchical.py:38-64): every feature of `meta["features"]`, cast to its dtype and
    reshaped to its `[frames, -1, width]` shape, in meta order."""
    meshes = {}
    for key, field in meta["features"].items():
        data = np.asarray(group[key][()]).astype(field["dtype"])
        meshes[key] = data.reshape(field["shape"])
    return meshes


def frame_as_mesh(traj: Mapping[str, np.ndarray], frame: int, meta: Dict[str, Any],
                  frame_target: Optional[int] = None):
    """`get_frame_as_mesh` (hierarchical.py:67-123): (mesh_pos, cells, point_data, target_point_data).  Point data =
    every feature but the topology keys at `frame`, then `node_type` of frame 0 LAST (that fixes the column order of
    x); targets = `meta["target_features"]` if given, else every dynamic feature, at `frame_target`."""
    target_point_data = None
    if frame_target is not None:
        names = meta.get("target_features")
        if names is None:
            names = [k for k, f in meta["features"].items() if f["type"] == "dynamic"]
        target_point_data = {k: traj[k][frame_target] for k in names}
    point_data = {k: traj[k][frame] for k in traj.keys() if k not in _TOPOLOGY_KEYS}
    point_data["node_type"] = traj["node_type"][0]
    mesh_pos = traj["mesh_pos"][frame] if traj["mesh_pos"].ndim > 1 else traj["mesh_pos"]
    cells = traj["cells"][frame] if traj["cells"].ndim > 1 else traj["cells"]
    return mesh_pos, cells, point_data, target_point_data


def _stack_features(arrays, extra=None) -> torch.Tensor:
    """Columns of a node-feature matrix (torch_graph.py:140-165): side by side if any array is 2-D, else each 1-D
    array is one column; computed in numpy's promoted dtype and rounded to fp32 once, as the reference does."""
    arrays = list(arrays)
    if any(a.ndim > 1 for a in arrays):
        cols = arrays + ([extra.reshape((-1, 1))] if extra is not None else [])
        mat = np.hstack(cols)
    else:
        mat = np.vstack(arrays + ([extra] if extra is not None else [])).T
    return torch.tensor(mat, dtype=torch.float32)


def faces_of_cells(cells: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """[C, 3] triangles or [C, 4] tetrahedra -> (face [3, F], tetra [4, C] or None); a tetrahedron gives its four
    triangles (v0v1v2, v1v2v3, v2v3v0, v3v0v1), grouped by kind (torch_graph.py:170-186)."""
    c = cells.t()
    if c.shape[0] == 4:
        face = torch.cat([c[0:3], c[1:4], torch.stack([c[2], c[3], c[0]]), torch.stack([c[3], c[0], c[1]])], dim=1)
        return face, c
    if c.shape[0] == 3:
        return c, None
    raise ValueError(f"cells must be triangles or tetrahedra, got {tuple(cells.shape)}")


def meshdata_to_graph(points: np.ndarray, cells: np.ndarray, point_data: Optional[Mapping[str, np.ndarray]],
                      time=1, target: Optional[Mapping[str, np.ndarray]] = None,
                      return_only_node_features: bool = False, id: Optional[str] = None):
    """`meshdata_to_graph` (torch_graph.py:114-193): x = [point data ... | time], y = target columns, pos, face, tetra."""
    if point_data is not None:
        x = _stack_features(point_data.values(), np.full((len(points),), time))
    else:
        x = torch.zeros((len(points), 1), dtype=torch.float32)
    if return_only_node_features:
        return x
    y = _stack_features(target.values()) if target is not None else None
    face, tetra = faces_of_cells(torch.as_tensor(np.asarray(cells)))
    return Graph(x=x, face=face, tetra=tetra, y=y, pos=torch.tensor(np.asarray(points), dtype=torch.float32), id=id)


def frame_as_graph(traj: Mapping[str, np.ndarray], frame: int, meta: Dict[str, Any],
                   frame_target: Optional[int] = None):
    """`get_frame_as_graph` (hierarchical.py:126-154): time = frame * meta.get("dt", 1)."""
    points, cells, point_data, target = frame_as_mesh(traj, frame, meta, frame_target)
    return meshdata_to_graph(points=points, cells=cells, point_data=point_data, time=frame * meta.get("dt", 1),
                             target=target)


def xdmf_frame_as_graph(reader, frame: int, meta: Dict[str, Any], target_step: int = 1, previous_step: Optional[int] = None,
                        mesh_id: Optional[str] = None):
    """The graph `XDMFDataset.__getitem__` builds from a time-series reader (xdmf_dataset.py:97-165) for given step
    offsets (the reference draws them at random in [1, random_next] / [1, random_prev]; here they are arguments):
    point data = the meta features present in the file, in META order, cast to the meta dtype, 1-D arrays as columns;
    targets = the dynamic ones at `frame + target_step`; x = [point data | time of the frame]; `target_dt` (and
    `previous_data`, `previous_dt` with `previous_step`) as the reference sets them.  `reader` = xdmf.XdmfTimeSeries or
    meshio's TimeSeriesReader."""
    num_steps = reader.num_steps
    if previous_step is not None and frame - previous_step < 0:
        previous_step = 1
    if frame + target_step > num_steps - 1:
        target_step = 1
    if frame >= num_steps - 1:
        raise IndexError(f"Frame index {frame} out of bounds for a trajectory with {num_steps} frames.")
    points, cells_dict = reader.read_points_cells()
    if not isinstance(cells_dict, dict):                       # meshio: a list of CellBlock
        cells_dict = {c.type: c.data for c in cells_dict}
    time, point_data, _ = reader.read_data(frame)
    _, target_point_data, _ = reader.read_data(frame + target_step)
    if "triangle" in cells_dict:
        cells = cells_dict["triangle"]
    elif "tetra" in cells_dict:
        cells = cells_dict["tetra"]
    else:
        raise ValueError("Unsupported cell type. Only 'triangle' and 'tetra' cells are supported.")
    feats = meta["features"]

    def pick(src, dynamic_only: bool):
        out = {}
        for k in feats:
            if k in src and (not dynamic_only or feats[k]["type"] == "dynamic"):
                v = np.array(src[k]).astype(feats[k]["dtype"])
                out[k] = v.reshape(-1, 1) if v.ndim == 1 else v
        return out

    graph = meshdata_to_graph(points=np.asarray(points).astype(np.float32), cells=np.asarray(cells),
                              point_data=pick(point_data, False), time=time, target=pick(target_point_data, True),
                              id=mesh_id)
    dt = meta.get("dt", 1) or 1
    graph.target_dt = target_step * dt
    if previous_step is not None:
        _, previous_data, _ = reader.read_data(frame - previous_step)
        graph.previous_data = pick(previous_data, True)
        graph.previous_dt = -previous_step * dt
    return graph


def masked_indexes(num_nodes: int, masking_ratio: float = 0.15, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """`get_masked_indexes` (utils/torch_graph.py:285-302): the first int((1 - ratio) n) entries of a host
    `randperm(n)` -- the indices of the nodes that are KEPT."""
    keep = int((1 - masking_ratio) * num_nodes)
    return torch.randperm(num_nodes, generator=generator)[:keep]


# ------------------------------------------------------------------------------------------------------- containers
class NpzContainer:
    """Trajectory groups in one `.npz`: arrays named `"<trajectory>/<feature>"` (what `save_npz_container` writes).
    `keys()` keeps the file's order, like `h5py.File.keys()` (get_h5_dataset, hierarchical.py:11-35)."""

    def __init__(self, path: str):
        self._z = np.load(path)
        self._index: "OrderedDict[str, Dict[str, str]]" = OrderedDict()
        for name in self._z.files:
            traj, _, feat = name.partition("/")
            if not feat:
                raise ValueError(f"{path}: array `{name}` is not `<trajectory>/<feature>`")
            self._index.setdefault(traj, {})[feat] = name

    def keys(self):
        return list(self._index.keys())

    def __getitem__(self, traj: str) -> Dict[str, np.ndarray]:
        return {feat: self._z[name] for feat, name in self._index[traj].items()}

    def close(self):
        self._z.close()


def save_npz_container(path: str, trajectories: Mapping[str, Mapping[str, np.ndarray]]) -> None:
    np.savez(path, **{f"{t}/{k}": np.asarray(v) for t, g in trajectories.items() for k, v in g.items()})


def open_container(path: str):
    """`.npz` -> NpzContainer; `.h5` / `.hdf5` -> `h5py.File(path, "r")`, or `h5lite.File` where h5py is not installed."""
    if path.endswith(".npz"):
        return NpzContainer(path)
    if path.endswith((".h5", ".hdf5")):
        try:
            import h5py
            return h5py.File(path, "r")
        except ImportError:
            from . import h5lite                # this package's read-only decoder (files written with default settings)
            return h5lite.File(path)
    raise ValueError(f"unknown trajectory container: {path}")


# ------------------------------------------------------------------------------------------------- resident on device
def trajectory_tensors(traj: Mapping[str, np.ndarray], meta: Dict[str, Any]):
    """All frames of a decoded trajectory at once: x [T, N, F] and y [T, N, D] as `frame_as_graph` builds them per
    frame (y[t] = the targets AT frame t: sample (t, t+1) reads y[t + 1]), pos [T or 1, N, dim], cells [T or 1, C, k].
    Host tensors; `ResidentTrajectory` moves them."""
    T = int(meta["trajectory_length"])
    dt = meta.get("dt", 1)
    names = meta.get("target_features")
    if names is None:
        names = [k for k, f in meta["features"].items() if f["type"] == "dynamic"]
    feats = [k for k in traj.keys() if k not in _TOPOLOGY_KEYS]
    n = traj["node_type"][0].shape[0]

    def per_frame(a: np.ndarray) -> np.ndarray:          # [T, N, w] whatever the stored frame count
        if a.shape[0] not in (1, T):
            raise ValueError(f"a feature has {a.shape[0]} frames, the trajectory {T}")
        return np.broadcast_to(a, (T,) + a.shape[1:])

    cols = [per_frame(traj[k]) for k in feats]
    cols.append(np.broadcast_to(traj["node_type"][0][None], (T, n) + traj["node_type"].shape[2:]))
    time = (np.arange(T) * dt).reshape(T, 1, 1)
    cols.append(np.broadcast_to(time, (T, n, 1)))
    x = torch.tensor(np.concatenate(cols, axis=2), dtype=torch.float32)
    y = torch.tensor(np.concatenate([per_frame(traj[k]) for k in names], axis=2), dtype=torch.float32)
    pos = torch.tensor(traj["mesh_pos"], dtype=torch.float32)
    cells = torch.as_tensor(traj["cells"])
    return x, y, pos, cells


class ResidentTrajectory:
    """One trajectory on the device: `x`, `y` ([T, N, .]), `pos`, and the topology run through the device set-up once.
    `moving_mesh` = positions or cells differ between frames; then edge_index / edge features are built per frame on
    demand (and cached), otherwise once."""

    def __init__(self, traj: Mapping[str, np.ndarray], meta: Dict[str, Any], device, add_edge_features: bool = True,
                 khop: int = 1):
        from . import _lib as L
        x, y, pos, cells = trajectory_tensors(traj, meta)
        pin = (lambda t: t.pin_memory()) if torch.cuda.is_available() else (lambda t: t)
        self.x = pin(x).to(device, non_blocking=True)
        self.y = pin(y).to(device, non_blocking=True)
        self.pos = pos.to(device)
        self.cells = cells.to(device)
        L.require_device(self.x)
        self.add_edge_features = add_edge_features
        self.khop = int(khop)
        self.num_frames, self.num_nodes = int(x.shape[0]), int(x.shape[1])
        same_cells = cells.shape[0] == 1 or bool((cells == cells[:1]).all())
        same_pos = pos.shape[0] == 1 or bool((pos == pos[:1]).all())
        self.moving_mesh = not (same_cells and same_pos)
        self._same_cells = same_cells
        self._topo: Dict[int, Tuple[torch.Tensor, Optional[torch.Tensor], torch.Tensor, Optional[torch.Tensor]]] = {}
        self.nbytes = sum(t.numel() * t.element_size() for t in (self.x, self.y, self.pos, self.cells))

    def topology(self, frame: int):
        """(edge_index, edge_attr or None, face, tetra) of `frame`; one entry serves every frame of a fixed mesh.  The
        SAME `edge_index` tensor object comes back every time, so `processor.csr_for` hits its cache."""
        from . import graph as G
        key = frame if self.moving_mesh else 0
        hit = self._topo.get(key)
        if hit is None:
            ci = key if self.cells.shape[0] > 1 else 0
            pi = key if self.pos.shape[0] > 1 else 0
            cells_key = 0 if self._same_cells else ci
            face, tetra = faces_of_cells(self.cells[ci])
            reuse = self._topo.get(("ei", cells_key))
            if reuse is None:
                ei = G.face_to_edge(face.contiguous(), self.num_nodes)
                if self.khop > 1:
                    ei = G.compute_k_hop_edge_index(ei, self.khop, self.num_nodes)
                self._topo[("ei", cells_key)] = ei
            else:
                ei = reuse
            ea = G.edge_features(self.pos[pi].contiguous(), ei) if self.add_edge_features else None
            hit = (ei, ea, face, tetra)
            self._topo[key] = hit
            self.nbytes += sum(t.numel() * t.element_size() for t in (ei, ea) if t is not None)
        return hit

    def sample(self, frame: int, frame_target: Optional[int], use_previous_data: bool = False):
        """The graph `H5Dataset.__getitem__` hands on after its preprocessing (x, y, pos, face, tetra, edge_index,
        edge_attr; `previous_data` when asked): views of the resident tensors."""
        if not 0 <= frame < self.num_frames or (frame_target is not None and not 0 <= frame_target < self.num_frames):
            raise IndexError(f"frame {frame} / target {frame_target} outside a {self.num_frames}-frame trajectory")
        ei, ea, face, tetra = self.topology(frame)
        g = Graph(x=self.x[frame], y=None if frame_target is None else self.y[frame_target],
                  pos=self.pos[frame if self.pos.shape[0] > 1 else 0], face=face, tetra=tetra, edge_index=ei,
                  edge_attr=ea)
        if use_previous_data:
            g.previous_data = self.x[frame - 1]
        return g


class TrajectoryDataset:
    """`H5Dataset` surface (h5_dataset.py:15-119, dataset.py:19-85) over resident trajectories:

        ds = TrajectoryDataset("train.npz", "meta.json", device="cuda")
        len(ds) == size_dataset * (trajectory_length - 1);   ds[i] -> Graph of (trajectory, frame), target = frame + 1

    `get_traj_frame` is the reference's index arithmetic, including its `use_previous_data` shift.  `graph.traj_index`
    is set as in the reference; `graph.csr` is the trajectory's cached receiver-sorted CSR (for `batch.collate`).
    `preprocessing` is the callable of `preprocessing.build_preprocessing` (noise, world edges, ...), applied as the
    reference applies its own; `new_edges_ratio` and `masking_ratio` act as in `BaseDataset` (dataset.py:104-137, 190-
    200: with a masking ratio an item is the pair (graph, kept node indices))."""

    def __init__(self, container_path: str, meta_path: str, device="cuda", preprocessing=None, khop: int = 1,
                 add_edge_features: bool = True, use_previous_data: bool = False, switch_to_val: bool = False,
                 resident_bytes: int = 64 << 30, masking_ratio: Optional[float] = None, new_edges_ratio: float = 0.0,
                 world_pos_index_start: Optional[int] = None, world_pos_index_end: Optional[int] = None):
        if switch_to_val:
            container_path = container_path.replace("train", "test")
        self.meta = load_meta(meta_path)
        self.trajectory_length = int(self.meta["trajectory_length"])
        self.dt = self.meta.get("dt", 1) or 1
        self.container = open_container(container_path)
        self.datasets_index = list(self.container.keys())
        self.device = torch.device(device)
        self.preprocessing = preprocessing
        self.khop = khop
        self.add_edge_features = add_edge_features
        self.use_previous_data = use_previous_data
        self.resident_bytes = int(resident_bytes)
        self.masking_ratio = masking_ratio
        self.new_edges_ratio = new_edges_ratio
        self.world_pos_index_start, self.world_pos_index_end = world_pos_index_start, world_pos_index_end
        # K-hop expansion: the reference applies it AFTER the preprocessing and caches the result per trajectory
        # (`_apply_k_hop`, dataset.py:139-175).  Without a preprocessing the expansion of the mesh edges is part of the
        # resident topology (same result, computed once); with one -- it may add world edges -- it runs here, on what
        # the preprocessing produced.
        self._khop_after = self.khop > 1 and preprocessing is not None
        self._khop_cache: Dict[int, Tuple[torch.Tensor, Optional[torch.Tensor]]] = {}
        self._resident: "OrderedDict[int, ResidentTrajectory]" = OrderedDict()

    @property
    def size_dataset(self) -> int:
        return len(self.datasets_index)

    def get_traj_frame(self, index: int) -> Tuple[int, int]:
        traj = index // (self.trajectory_length - 1)
        frame = index % (self.trajectory_length - 1) + int(self.use_previous_data)
        return traj, frame

    def __len__(self) -> int:
        return self.size_dataset * (self.trajectory_length - 1)

    def trajectory(self, traj_index: int) -> ResidentTrajectory:
        hit = self._resident.get(traj_index)
        if hit is not None:
            self._resident.move_to_end(traj_index)
            return hit
        group = self.container[self.datasets_index[traj_index]]
        rt = ResidentTrajectory(traj_as_meshes(group, self.meta), self.meta, self.device,
                                add_edge_features=self.add_edge_features, khop=1 if self._khop_after else self.khop)
        self._resident[traj_index] = rt
        while len(self._resident) > 1 and sum(t.nbytes for t in self._resident.values()) > self.resident_bytes:
            self._resident.popitem(last=False)
        return rt

    def __getitem__(self, index: int):
        if not 0 <= index < len(self):
            raise IndexError(index)
        traj_index, frame = self.get_traj_frame(index)
        rt = self.trajectory(traj_index)
        g = rt.sample(frame, frame + 1, use_previous_data=self.use_previous_data)
        if self.preprocessing is not None:
            g = self.preprocessing(g)
        if self._khop_after:
            g = self._apply_k_hop(g, traj_index, rt.num_nodes)
        if 0.0 < self.new_edges_ratio <= 1.0:
            g = self._add_random_edges(g, rt.num_nodes)
        if hasattr(g, "previous_data"):
            del g.previous_data
        g.traj_index = traj_index
        from . import processor as _proc
        g.csr = _proc.csr_for(g.edge_index, rt.num_nodes)
        if self.masking_ratio is not None:
            return g, masked_indexes(rt.num_nodes, self.masking_ratio)
        return g

    def _apply_k_hop(self, g, traj_index: int, num_nodes: int):
        """`BaseDataset._apply_k_hop` (dataset.py:139-175) + `compute_k_hop_graph` (utils/torch_graph.py:56-112): the
        K-hop edge set of the PREPROCESSED graph, with Cartesian / Distance features recomputed on it (and the world-
        position features when the index range is configured); computed for the first sample of a trajectory and
        reused for its other frames, as the reference's per-trajectory cache does."""
        from . import graph as G
        hit = self._khop_cache.get(traj_index)
        if hit is None:
            ei = G.compute_k_hop_edge_index(g.edge_index, self.khop, num_nodes)
            ea = None
            if self.add_edge_features:
                ea = G.edge_features(g.pos.contiguous(), ei)
                if self.world_pos_index_start is not None and self.world_pos_index_end is not None:
                    wp = g.x[:, self.world_pos_index_start:self.world_pos_index_end].contiguous()
                    ea = G.world_pos_features(ea, wp, ei)
            hit = (ei, ea)
            self._khop_cache[traj_index] = hit
        g.edge_index = hit[0]
        if self.add_edge_features:
            g.edge_attr = hit[1]
        return g

    def _add_random_edges(self, g, num_nodes: int):
        """`BaseDataset._add_random_edges` (dataset.py:104-137): round(E p) // 2 random undirected pairs appended, edge
        features recomputed on the new edge set (a per-sample edge_index: its CSR is sorted per sample)."""
        from . import graph as G
        g.edge_index, _ = G.add_random_edges(g.edge_index, self.new_edges_ratio, num_nodes)
        if self.add_edge_features:
            g.edge_attr = G.edge_features(g.pos.contiguous(), g.edge_index)
        return g

    def close(self):
        if getattr(self, "container", None) is not None:
            self.container.close()
            self.container = None
