"""ORACLE / TEST INFRASTRUCTURE ONLY -- stdlib reader for the reference's VTU fixtures.

The reference reads its mock meshes with meshio (`tests/mock.py:21-27`,
`graphphysics/dataset/xdmf_dataset.py:70-186`); meshio is not installed here, so
this file decodes the same files (VTK XML UnstructuredGrid, "binary" = base64 of
a uint32 block header followed by zlib blocks) with base64 + zlib only.  It is
used by `tests/golden/make_golden.py` to turn `tests/mock_vtu*/*.vtu` into the
small `.npz` fixtures committed under `tests/golden/`.  Nothing in the product
path imports it.
"""
import base64
import re
import struct
import zlib

import numpy as np

_DTYPES = {
    "Float32": np.float32, "Float64": np.float64,
    "Int32": np.int32, "Int64": np.int64, "UInt8": np.uint8,
    "UInt32": np.uint32, "UInt64": np.uint64, "Int8": np.int8,
}


def _decode_block(text: str, dtype) -> np.ndarray:
    """vtkZLibDataCompressor payload: header = [nblocks, blocksize, lastsize, csize...]
    as uint32, base64-encoded on its own; compressed blocks follow, base64-encoded
    together."""
    text = text.strip()
    # header is 3 + nblocks uint32; decode the first word to learn nblocks
    first = base64.b64decode(text[:8])  # 6 bytes -> first uint32 is valid
    nblocks = struct.unpack("<I", first[:4])[0]
    hdr_bytes = 4 * (3 + nblocks)
    hdr_b64 = 4 * ((hdr_bytes + 2) // 3)
    header = np.frombuffer(base64.b64decode(text[:hdr_b64])[:hdr_bytes], dtype="<u4")
    csizes = header[3:]
    payload = base64.b64decode(text[hdr_b64:])
    out = bytearray()
    off = 0
    for cs in csizes:
        out += zlib.decompress(payload[off:off + int(cs)])
        off += int(cs)
    return np.frombuffer(bytes(out), dtype=np.dtype(dtype).newbyteorder("<")).copy()


def read_vtu(path: str):
    """Returns (points [N,3] f32, cells [C, nodes_per_cell] int64, point_data dict)."""
    with open(path, "r") as f:
        xml = f.read()
    piece = re.search(r'<Piece NumberOfPoints="(\d+)" NumberOfCells="(\d+)"', xml)
    n_points, n_cells = int(piece.group(1)), int(piece.group(2))
    arrays = {}
    sections = {}
    for sec in ("Points", "Cells", "PointData", "CellData"):
        m = re.search(r"<%s[^>]*>(.*?)</%s>" % (sec, sec), xml, flags=re.S)
        sections[sec] = m.group(1) if m else ""
    out = {}
    for sec, body in sections.items():
        d = {}
        for m in re.finditer(r"<DataArray([^>]*)>(.*?)</DataArray>", body, flags=re.S):
            attrs = dict(re.findall(r'(\w+)="([^"]*)"', m.group(1)))
            arr = _decode_block(m.group(2), _DTYPES[attrs["type"]])
            nc = int(attrs.get("NumberOfComponents", "1"))
            if nc > 1:
                arr = arr.reshape(-1, nc)
            d[attrs.get("Name", "")] = arr
        out[sec] = d
    points = out["Points"]["Points"].astype(np.float32)
    assert points.shape[0] == n_points
    conn = out["Cells"]["connectivity"].astype(np.int64)
    offsets = out["Cells"]["offsets"].astype(np.int64)
    assert offsets.shape[0] == n_cells
    per = np.diff(np.concatenate([[0], offsets]))
    assert (per == per[0]).all(), "mixed cell types not supported by this reader"
    cells = conn.reshape(n_cells, int(per[0]))
    return points, cells, out["PointData"]
