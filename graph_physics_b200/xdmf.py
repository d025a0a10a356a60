"""XDMF time series in: the reader side of `meshio.xdmf.TimeSeriesReader` as the reference uses it
(`graphphysics/dataset/xdmf_dataset.py:97-118`: `num_steps`, `read_points_cells()`, `read_data(k)`), for the files
meshio's `TimeSeriesWriter` produces (`utils/meshio_mesh.py:70-113`, `training/lightning_module.py:124-161`): a temporal
collection of grids that include one shared mesh grid, heavy data in an HDF5 file next to the `.xdmf` (`Format="HDF"`,
`file.h5:/dataN`) or inline (`Format="XML"`).  A single uniform grid (one frame, e.g. tests/mock_airfoil) reads as a
series of one step.  Heavy data is read with h5py when it is installed, else with `h5lite` (this package's decoder).

Only what the path needs: node-centred and cell-centred scalar / vector attributes, one topology per file.
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
