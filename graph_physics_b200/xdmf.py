"""XDMF time series in: the reader side of `meshio.xdmf.TimeSeriesReader` as the reference uses it
(`graphphysics/dataset/xdmf_dataset.py:97-118`: `num_steps`, `read_points_cells()`, `read_data(k)`), for the files
meshio's `TimeSeriesWriter` produces (`utils/meshio_mesh.py:70-113`, `training/lightning_module.py:124-161`): a temporal
collection of grids that include one shared mesh grid, heavy data in an HDF5 file next to the `.xdmf` (`Format="HDF"`,
`file.h5:/dataN`), in a raw binary file (`Format="Binary"`, `Seek`, `Endian`) or inline (`Format="XML"`).  A single uniform grid (one frame, e.g. tests/mock_airfoil) reads as a
series of one step.  Heavy data is read with h5py when it is installed, else with `h5lite` (this package's decoder).

Only what the path needs: node-centred and cell-centred scalar / vector attributes, one topology per file.

`XdmfTimeSeriesWriter` is the writer side (`meshio.xdmf.TimeSeriesWriter` as used by `vtu_to_xdmf`,
meshio_mesh.py:70-113, and `_save_trajectory_to_xdmf`, lightning_module.py:124-161: `write_points_cells` once, then
`write_data(t, point_data=...)` per frame) with the same XML structure -- a temporal collection whose grids include the
mesh grid by XPointer -- and the heavy data in ONE raw little-endian file (`Format="Binary"`, which ParaView and meshio
read) instead of HDF5, which cannot be written here without libhdf5.
"""
from __future__ import annotations

import os
import xml.etree.ElementTree as ET
from typing import Dict, List, Tuple

import numpy as np

# XDMF topology name -> (meshio cell type, nodes per cell)
_TOPOLOGY = {"triangle": ("triangle", 3), "tetrahedron": ("tetra", 4), "quadrilateral": ("quad", 4),
             "polyline": ("line", 2), "hexahedron": ("hexahedron", 8)}
_NP = {("Float", "4"): np.float32, ("Float", "8"): np.float64, ("Int", "4"): np.int32, ("Int", "8"): np.int64,
       ("UInt", "4"): np.uint32, ("UInt", "8"): np.uint64, ("Int", "1"): np.int8, ("UInt", "1"): np.uint8}


def _open_h5(path: str):
    try:
        import h5py
        return h5py.File(path, "r")
    except ImportError:
        from . import h5lite
        return h5lite.File(path)


class XdmfTimeSeries:
    """with XdmfTimeSeries(path) as reader: reader.num_steps; reader.read_points_cells(); reader.read_data(k)"""

    def __init__(self, filename: str):
        self.filename = filename
        self._dir = os.path.dirname(os.path.abspath(filename))
        self._h5: Dict[str, object] = {}
        root = ET.parse(filename).getroot()
        if root.tag != "Xdmf":
            raise ValueError(f"{filename}: not an XDMF file")
        domain = root.find("Domain")
        grids = domain.findall("Grid") if domain is not None else []
        self._mesh = None
        self._steps: List[ET.Element] = []
        for g in grids:
            if g.get("GridType") == "Collection" and g.get("CollectionType") == "Temporal":
                self._steps = g.findall("Grid")
            elif self._mesh is None and g.find("Topology") is not None:
                self._mesh = g
        if self._mesh is None:
            raise ValueError(f"{filename}: no grid with a topology")
        if not self._steps:                      # a single uniform grid: one step that carries its own attributes
            self._steps = [self._mesh]
        self.num_steps = len(self._steps)

    # ---- heavy data
    def _item(self, item: ET.Element) -> np.ndarray:
        dims = [int(d) for d in item.get("Dimensions", "").split()]
        fmt = item.get("Format", "XML")
        dtype = _NP.get((item.get("DataType", item.get("NumberType", "Float")), item.get("Precision", "4")))
        if dtype is None:
            raise ValueError(f"{self.filename}: unsupported DataItem type {item.attrib}")
        text = (item.text or "").strip()
        if fmt == "XML":
            return np.array(text.split(), dtype=dtype).reshape(dims)
        if fmt == "Binary":                      # raw file next to the .xdmf: `Seek` bytes in, `Endian` byte order
            order = {"Little": "<", "Big": ">", "Native": "="}[item.get("Endian", "Native")]
            n = int(np.prod(dims)) if dims else 1
            arr = np.fromfile(os.path.join(self._dir, text), dtype=np.dtype(dtype).newbyteorder(order), count=n,
                              offset=int(item.get("Seek", "0")))
            if arr.size != n:
                raise ValueError(f"{self.filename}: {text} is shorter than the DataItem says")
            return arr.astype(dtype).reshape(dims)
        if fmt != "HDF":
            raise ValueError(f"{self.filename}: unsupported DataItem format {fmt}")
        fname, _, key = text.partition(":")
        handle = self._h5.get(fname)
        if handle is None:
            handle = self._h5[fname] = _open_h5(os.path.join(self._dir, fname))
        arr = np.asarray(handle[key.lstrip("/")][()])
        if list(arr.shape) != dims:
            raise ValueError(f"{self.filename}: {text} has shape {arr.shape}, the DataItem says {dims}")
        return arr

    # ---- meshio.xdmf.TimeSeriesReader surface
    def read_points_cells(self) -> Tuple[np.ndarray, Dict[str, np.ndarray]]:
        geo, topo = self._mesh.find("Geometry"), self._mesh.find("Topology")
        points = self._item(geo.find("DataItem"))
        name = topo.get("TopologyType", topo.get("Type", "")).lower()
        if name not in _TOPOLOGY:
            raise ValueError(f"{self.filename}: unsupported topology {name}")
        ctype, k = _TOPOLOGY[name]
        cells = self._item(topo.find("DataItem")).reshape(-1, k)
        return points, {ctype: cells}

    def read_data(self, k: int):
        """(time, point_data, cell_data) of step k."""
        if not 0 <= k < self.num_steps:
            raise IndexError(k)
        grid = self._steps[k]
        t = grid.find("Time")
        time = float(t.get("Value")) if t is not None else 0.0
        point_data: Dict[str, np.ndarray] = {}
        cell_data: Dict[str, np.ndarray] = {}
        for a in grid.findall("Attribute"):
            arr = self._item(a.find("DataItem"))
            (point_data if a.get("Center", "Node") == "Node" else cell_data)[a.get("Name")] = arr
        return time, point_data, cell_data

    def close(self):
        for h in self._h5.values():
            h.close()
        self._h5 = {}

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


_XDMF_TYPE = {np.dtype("float32"): ("Float", 4), np.dtype("float64"): ("Float", 8), np.dtype("int32"): ("Int", 4),
              np.dtype("int64"): ("Int", 8), np.dtype("uint8"): ("UInt", 1), np.dtype("int8"): ("Int", 1),
              np.dtype("uint32"): ("UInt", 4), np.dtype("uint64"): ("UInt", 8)}
_TOPOLOGY_OUT = {"triangle": "Triangle", "tetra": "Tetrahedron", "quad": "Quadrilateral", "line": "Polyline",
                 "hexahedron": "Hexahedron"}


class XdmfTimeSeriesWriter:
    """with XdmfTimeSeriesWriter("out.xdmf") as w: w.write_points_cells(points, {"triangle": cells});
    w.write_data(t, point_data={...})  -> out.xdmf + out.bin"""

    def __init__(self, filename: str):
        if not filename.endswith(".xdmf"):
            raise ValueError("XdmfTimeSeriesWriter: the file name must end in .xdmf")
        self.filename = filename
        self._bin_name = os.path.basename(filename)[:-len(".xdmf")] + ".bin"
        if os.path.dirname(filename):
            os.makedirs(os.path.dirname(filename), exist_ok=True)
        self._bin = open(os.path.join(os.path.dirname(os.path.abspath(filename)), self._bin_name), "wb")
        self._root = ET.Element("Xdmf", Version="3.0")
        self._root.set("xmlns:xi", "http://www.w3.org/2001/XInclude")
        domain = ET.SubElement(self._root, "Domain")
        self._collection = ET.SubElement(domain, "Grid", Name="TimeSeries_b200", GridType="Collection",
                                         CollectionType="Temporal")
        self._domain = domain
        self._has_mesh = False
        self._num_points = self._num_cells = 0

    def _item(self, parent: ET.Element, a: np.ndarray) -> None:
        a = np.ascontiguousarray(a)
        if a.dtype not in _XDMF_TYPE:
            a = a.astype(np.float64 if a.dtype.kind == "f" else np.int64)
        dtype, prec = _XDMF_TYPE[a.dtype]
        item = ET.SubElement(parent, "DataItem", DataType=dtype, Dimensions=" ".join(str(d) for d in a.shape),
                             Format="Binary", Precision=str(prec), Endian="Little", Seek=str(self._bin.tell()))
        item.text = self._bin_name
        self._bin.write(a.astype(a.dtype.newbyteorder("<"), copy=False).tobytes())

    def write_points_cells(self, points: np.ndarray, cells) -> None:
        if self._has_mesh:
            raise ValueError("XdmfTimeSeriesWriter: the mesh is written once")
        if not isinstance(cells, dict):                     # meshio style: [(type, array)] or CellBlocks
            cells = {getattr(c, "type", c[0]): getattr(c, "data", c[1]) for c in cells}
        if len(cells) != 1:
            raise ValueError("XdmfTimeSeriesWriter: one cell type per file")
        (ctype, data), = cells.items()
        if ctype not in _TOPOLOGY_OUT:
            raise ValueError(f"XdmfTimeSeriesWriter: unsupported cell type {ctype}")
        points, data = np.asarray(points), np.asarray(data)
        grid = ET.SubElement(self._domain, "Grid", Name="mesh", GridType="Uniform")
        geo = ET.SubElement(grid, "Geometry", GeometryType="XY" if points.shape[1] == 2 else "XYZ")
        self._item(geo, points)
        topo = ET.SubElement(grid, "Topology", TopologyType=_TOPOLOGY_OUT[ctype], NumberOfElements=str(data.shape[0]))
        self._item(topo, data)
        self._has_mesh = True
        self._num_points, self._num_cells = points.shape[0], data.shape[0]

    def write_data(self, t: float, point_data=None, cell_data=None) -> None:
        if not self._has_mesh:
            raise ValueError("XdmfTimeSeriesWriter: write_points_cells first")
        for center, data, n in (("node", point_data or {}, self._num_points), ("cell", cell_data or {}, self._num_cells)):
            for name, a in data.items():        # validate everything before the step is added to the document
                a = np.asarray(a)
                if a.ndim not in (1, 2) or a.shape[0] != n:
                    raise ValueError(f"XdmfTimeSeriesWriter: `{name}` has shape {a.shape} for {n} {center}s")
        grid = ET.SubElement(self._collection, "Grid")
        inc = ET.SubElement(grid, "xi:include")
        inc.set("xpointer", 'xpointer(//Grid[@Name="mesh"]/*[self::Topology or self::Geometry])')
        ET.SubElement(grid, "Time", Value=repr(float(t)) if float(t) != int(t) else str(int(t)))
        for center, data, n in (("Node", point_data or {}, self._num_points), ("Cell", cell_data or {}, self._num_cells)):
            for name, a in data.items():
                a = np.asarray(a)
                kind = "Scalar" if a.ndim == 1 or a.shape[1] == 1 else "Vector"
                att = ET.SubElement(grid, "Attribute", Name=name, AttributeType=kind, Center=center)
                self._item(att, a)

    def close(self) -> None:
        if self._bin is not None:
            self._bin.close()
            self._bin = None
            ET.ElementTree(self._root).write(self.filename)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def write_trajectory(save_dir: str, archive_filename: str, trajectory, timestep: float = 1.0,
                     use_previous_data: bool = False) -> str:
    """`_save_trajectory_to_xdmf` (lightning_module.py:124-161): the mesh of the first graph once, then all node data
    (`x0..`, `y0..`, as `convert_to_meshio_vtu(add_all_data=True)` names them) of every graph at times `timestep`,
    `2 timestep`, ... (starting at `2 timestep` with previous data).  Returns the `.xdmf` path."""
    from .vtu import mesh_arrays
    path = os.path.join(save_dir, f"{archive_filename}.xdmf")
    t = timestep if not use_previous_data else 2 * timestep
    with XdmfTimeSeriesWriter(path) as w:
        for i, graph in enumerate(trajectory):
            points, cells, point_data = mesh_arrays(graph, add_all_data=True)
            if i == 0:
                w.write_points_cells(points, {"tetra" if cells.shape[1] == 4 else "triangle": cells})
            w.write_data(t, point_data=point_data)
            t += timestep
    return path
