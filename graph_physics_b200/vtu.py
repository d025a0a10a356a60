"""Mesh files either side of the path: VTK XML UnstructuredGrid (`.vtu`) in and out, plus a
`.pvd` time-series index -- the role meshio plays in the reference (`graphphysics/utils/meshio_mesh.py:10-67`,
`convert_to_meshio_vtu`; `dataset/xdmf_dataset.py` / `tests/mock.py:21-27` read meshes through `meshio.read`;
`training/lightning_module.py:124-161` writes a predicted trajectory as a time series).  meshio is not installed and
XDMF keeps its heavy data in HDF5 (no h5py here), so predicted trajectories are written as one `.vtu` per frame and a
`.pvd` collection, which ParaView and `meshio.read` open directly.

`write_vtu(path, graph, add_all_data=True)` follows `convert_to_meshio_vtu`: positions zero-padded to three coordinates
(more than three raises), cells = `graph.tetra` when present else `graph.face` (stored [k, C], written as rows), point
data `x0..x{F-1}` from `graph.x` and `y0..` from `graph.y`.  The encoding is what meshio itself writes by default and
what the reference's fixtures use: `format="binary"`, zlib blocks (`vtkZLibDataCompressor`), uint32 block header,
little endian.  `read_vtu` decodes that encoding and plain `ascii` arrays, for files with one cell type.

Host-side, stdlib + numpy: tensors are brought to the host once per frame (this is output, off the step path).
"""
from __future__ import annotations

import base64
import os
import re
import struct
import zlib
from typing import Dict, Mapping, Optional, Sequence, Tuple

import numpy as np
import torch

_VTK_TYPE = {np.dtype("float32"): "Float32", np.dtype("float64"): "Float64", np.dtype("int32"): "Int32",
             np.dtype("int64"): "Int64", np.dtype("uint8"): "UInt8", np.dtype("int8"): "Int8",
             np.dtype("uint32"): "UInt32", np.dtype("uint64"): "UInt64"}
_NP_TYPE = {v: k for k, v in _VTK_TYPE.items()}
# VTK cell type ids by nodes per cell for the simplices this package handles (VTK_TRIANGLE = 5, VTK_TETRA = 10)
_CELL_TYPE = {3: 5, 4: 10}
_BLOCK = 32768


def _encode(a: np.ndarray) -> str:
    raw = np.ascontiguousarray(a).astype(a.dtype.newbyteorder("<"), copy=False).tobytes()
    blocks = [raw[i:i + _BLOCK] for i in range(0, len(raw), _BLOCK)] or [b""]
    comp = [zlib.compress(b) for b in blocks]
    last = len(blocks[-1]) if len(raw) % _BLOCK or not raw else _BLOCK
    header = struct.pack("<%dI" % (3 + len(comp)), len(blocks), _BLOCK, last, *[len(c) for c in comp])
    return base64.b64encode(header).decode() + base64.b64encode(b"".join(comp)).decode()


def _decode(text: str, dtype: np.dtype) -> np.ndarray:
    text = text.strip()
    nblocks = struct.unpack("<I", base64.b64decode(text[:8])[:4])[0]
    hdr_bytes = 4 * (3 + nblocks)
    hdr_b64 = 4 * ((hdr_bytes + 2) // 3)
    header = np.frombuffer(base64.b64decode(text[:hdr_b64])[:hdr_bytes], dtype="<u4")
    payload = base64.b64decode(text[hdr_b64:])
    out, off = bytearray(), 0
    for cs in header[3:]:
        out += zlib.decompress(payload[off:off + int(cs)])
        off += int(cs)
    return np.frombuffer(bytes(out), dtype=dtype.newbyteorder("<")).astype(dtype)


def _data_array(name: str, a: np.ndarray, components: int = 1) -> str:
    comp = f' NumberOfComponents="{components}"' if components > 1 else ""
    return (f'<DataArray type="{_VTK_TYPE[a.dtype]}" Name="{name}"{comp} format="binary">\n{_encode(a)}\n</DataArray>\n')


def _host(t) -> np.ndarray:
    return t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t)


def mesh_arrays(graph, add_all_data: bool = False) -> Tuple[np.ndarray, np.ndarray, Dict[str, np.ndarray]]:
    """(points [N, 3], cells [C, k], point_data) of a graph, as `convert_to_meshio_vtu` (meshio_mesh.py:10-67) builds
    its `meshio.Mesh`."""
    if getattr(graph, "pos", None) is None:
        raise ValueError("Graph must have 'pos' attribute with node positions.")
    vertices = _host(graph.pos)
    dim = vertices.shape[1]
    if dim > 3:
        raise ValueError(f"Unsupported vertex dimension: {dim}")
    if dim < 3:
        vertices = np.hstack([vertices, np.zeros((vertices.shape[0], 3 - dim), dtype=vertices.dtype)])
    if getattr(graph, "face", None) is None:
        raise ValueError("Graph must have 'face' attribute with face indices.")
    tetra = getattr(graph, "tetra", None)
    cells = _host(tetra if tetra is not None else graph.face).T
    point_data: Dict[str, np.ndarray] = {}
    if add_all_data and getattr(graph, "x", None) is not None:
        x = _host(graph.x)
        for i in range(x.shape[1]):
            point_data[f"x{i}"] = x[:, i]
    if add_all_data and getattr(graph, "y", None) is not None:
        y = _host(graph.y)
        for i in range(y.shape[1]):
            point_data[f"y{i}"] = y[:, i]
    return vertices, cells, point_data


def write_vtu_arrays(path: str, points: np.ndarray, cells: np.ndarray,
                     point_data: Optional[Mapping[str, np.ndarray]] = None) -> None:
    points = np.ascontiguousarray(points)
    cells = np.ascontiguousarray(cells)
    n, (c, k) = points.shape[0], cells.shape
    if points.shape[1] != 3 or k not in _CELL_TYPE:
        raise ValueError(f"write_vtu: points must be [N, 3] and cells triangles or tetrahedra, got {points.shape}, {cells.shape}")
    if points.dtype not in _VTK_TYPE:
        points = points.astype(np.float32)
    conn = cells.reshape(-1).astype(np.int64)
    offsets = (np.arange(1, c + 1, dtype=np.int64) * k)
    types = np.full(c, _CELL_TYPE[k], dtype=np.uint8)
    parts = ['<?xml version="1.0"?>\n<VTKFile type="UnstructuredGrid" version="0.1" byte_order="LittleEndian" '
             'compressor="vtkZLibDataCompressor">\n<UnstructuredGrid>\n',
             f'<Piece NumberOfPoints="{n}" NumberOfCells="{c}">\n<Points>\n', _data_array("Points", points, 3),
             "</Points>\n<Cells>\n", _data_array("connectivity", conn), _data_array("offsets", offsets),
             _data_array("types", types), "</Cells>\n"]
    if point_data:
        parts.append("<PointData>\n")
        for name, a in point_data.items():
            a = np.ascontiguousarray(a)
            if a.shape[0] != n or a.ndim > 2:
                raise ValueError(f"write_vtu: point data `{name}` has shape {a.shape} for {n} points")
            if a.dtype not in _VTK_TYPE:
                a = a.astype(np.float64 if a.dtype.kind == "f" else np.int64)
            parts.append(_data_array(name, a, a.shape[1] if a.ndim == 2 else 1))
        parts.append("</PointData>\n")
    parts.append("</Piece>\n</UnstructuredGrid>\n</VTKFile>\n")
    if os.path.dirname(path):
        os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write("".join(parts))


def write_vtu(path: str, graph, add_all_data: bool = False) -> None:
    """`convert_to_meshio_vtu(graph, add_all_data)` followed by `mesh.write(path)`."""
    write_vtu_arrays(path, *mesh_arrays(graph, add_all_data))


def read_vtu(path: str) -> Tuple[np.ndarray, np.ndarray, Dict[str, np.ndarray]]:
    """(points [N, 3], cells [C, k] int64, point_data) of a one-cell-type `.vtu` (zlib-binary or ascii arrays)."""
    with open(path, "r") as f:
        xml = f.read()
    piece = re.search(r'<Piece NumberOfPoints="(\d+)"\s+NumberOfCells="(\d+)"', xml)
    if piece is None:
        raise ValueError(f"{path}: not a VTK UnstructuredGrid piece")
    n_points, n_cells = int(piece.group(1)), int(piece.group(2))
    compressed = "vtkZLibDataCompressor" in xml[:xml.index("<UnstructuredGrid")]
    out: Dict[str, Dict[str, np.ndarray]] = {}
    for sec in ("Points", "Cells", "PointData"):
        m = re.search(r"<%s[^>]*>(.*?)</%s>" % (sec, sec), xml, flags=re.S)
        d: Dict[str, np.ndarray] = {}
        for a in re.finditer(r"<DataArray([^>]*)>(.*?)</DataArray>", m.group(1) if m else "", flags=re.S):
            attrs = dict(re.findall(r'(\w+)="([^"]*)"', a.group(1)))
            dtype = _NP_TYPE[attrs["type"]]
            fmt = attrs.get("format", "ascii")
            if fmt == "ascii":
                arr = np.array(a.group(2).split(), dtype=dtype)
            elif fmt == "binary" and compressed:
                arr = _decode(a.group(2), dtype)
            else:
                raise ValueError(f"{path}: unsupported DataArray encoding `{fmt}` (compressed={compressed})")
            nc = int(attrs.get("NumberOfComponents", "1"))
            d[attrs.get("Name", "")] = arr.reshape(-1, nc) if nc > 1 else arr
        out[sec] = d
    points = out["Points"]["Points"]
    conn = out["Cells"]["connectivity"].astype(np.int64)
    offsets = out["Cells"]["offsets"].astype(np.int64)
    if points.shape[0] != n_points or offsets.shape[0] != n_cells:
        raise ValueError(f"{path}: array sizes do not match the piece header")
    per = np.diff(np.concatenate([[0], offsets]))
    if n_cells and not (per == per[0]).all():
        raise ValueError(f"{path}: mixed cell types are not supported")
    cells = conn.reshape(n_cells, int(per[0]) if n_cells else 3)
    return points, cells, out["PointData"]


def write_trajectory(save_dir: str, archive_filename: str, trajectory: Sequence, timestep: float = 1.0,
                     start_time: Optional[float] = None) -> str:
    """A predicted trajectory (`_save_trajectory_to_xdmf`, lightning_module.py:124-161: all node data of every frame,
    times `timestep`, `2 * timestep`, ...) as `<archive>_<i>.vtu` files plus `<archive>.pvd`; returns the `.pvd` path."""
    os.makedirs(save_dir, exist_ok=True)
    t = timestep if start_time is None else start_time
    entries = []
    for i, graph in enumerate(trajectory):
        name = f"{archive_filename}_{i:05d}.vtu"
        write_vtu(os.path.join(save_dir, name), graph, add_all_data=True)
        entries.append(f'<DataSet timestep="{t:.9g}" part="0" file="{name}"/>\n')
        t += timestep
    pvd = os.path.join(save_dir, f"{archive_filename}.pvd")
    with open(pvd, "w") as f:
        f.write('<?xml version="1.0"?>\n<VTKFile type="Collection" version="0.1" byte_order="LittleEndian">\n'
                "<Collection>\n" + "".join(entries) + "</Collection>\n</VTKFile>\n")
    return pvd
