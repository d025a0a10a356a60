"""A small read-only HDF5 decoder: enough of the file format to open the trajectory archives the
reference reads with h5py (`graphphysics/utils/hierarchical.py:11-64`: `File[traj][feature][()]`) and the heavy-data
files next to its XDMF time series (`dataset/xdmf_dataset.py:70-186`, written by meshio) in an environment without
h5py / libhdf5.  stdlib + numpy.

Covered -- what h5py and meshio write with default settings (`libver="earliest"`):
  * superblock versions 0 and 1, 8-byte (or 4-byte) offsets / lengths;
  * old-style groups: symbol-table message -> version-1 B-tree of symbol-table nodes + local heap (nested groups too);
  * version-1 object headers with continuation blocks; dataspace (v1, v2), datatype (fixed-point and IEEE floating
    point, either byte order), data layout v3 (compact, contiguous, chunked through a version-1 chunk B-tree), filter
    pipeline v1 / v2 with deflate, shuffle and fletcher32;
  * `File.keys()`, `File[name]` / `Group[name]` (paths with `/` too), `Dataset.shape`, `.dtype`, `Dataset[()]`.
Not covered (raises `H5Unsupported`, never guesses): new-style groups / version-2 object headers (files written with
`libver="latest"`), variable-length or compound types, external storage, other filters (szip, lzf, ...).

The interface is the subset of h5py the reference uses, so `dataset.open_container("x.h5")` hands either this or a
real `h5py.File` to the same code.
"""
from __future__ import annotations

import struct
import zlib
from typing import Dict, List, Optional, Tuple

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Unsupported(NotImplementedError):
    pass


class _Reader:
    def __init__(self, buf: bytes):
        self.b = buf
        self.O = 8      # size of offsets
        self.L = 8      # size of lengths

    def u(self, off: int, n: int) -> int:
        return int.from_bytes(self.b[off:off + n], "little")

    def addr(self, off: int) -> int:
        v = self.u(off, self.O)
        return _UNDEF if v == (1 << (8 * self.O)) - 1 else v


class Dataset:
    def __init__(self, f: "File", name: str, msgs: List[Tuple[int, int, int]]):
        self._f, self.name = f, name
        r = f._r
        self.shape: Tuple[int, ...] = ()
        self.dtype: Optional[np.dtype] = None
        self._layout = None
        self._filters: List[Tuple[int, List[int]]] = []
        for mtype, off, size in msgs:
            if mtype == 0x1:
                self.shape = _dataspace(r, off)
            elif mtype == 0x3:
                self.dtype = _datatype(r, off)
            elif mtype == 0x8:
                self._layout = _layout(r, off)
            elif mtype == 0xB:
                self._filters = _filters(r, off)
        if self.dtype is None or self._layout is None:
            raise H5Unsupported(f"{name}: object is not a simple dataset")

    def __getitem__(self, key):
        arr = self._read()
        return arr if key == () or key is Ellipsis else arr[key]

    def _read(self) -> np.ndarray:
        r, (kind, info) = self._f._r, self._layout
        n = int(np.prod(self.shape, dtype=np.int64)) if self.shape else 1
        nbytes = n * self.dtype.itemsize
        if kind == "compact":
            raw = info
        elif kind == "contiguous":
            address, _size = info
            raw = bytes(nbytes) if address == _UNDEF else r.b[address:address + nbytes]
        else:
            return self._read_chunked(info)
        if len(raw) < nbytes:
            raise ValueError(f"{self.name}: truncated data")
        return np.frombuffer(raw[:nbytes], dtype=self.dtype).reshape(self.shape).astype(self.dtype.newbyteorder("="))

    def _read_chunked(self, info) -> np.ndarray:
        btree, chunk_dims = info
        rank = len(self.shape)
        if len(chunk_dims) != rank + 1 or chunk_dims[-1] != self.dtype.itemsize:
            raise H5Unsupported(f"{self.name}: unexpected chunk description {chunk_dims}")
        cshape = tuple(chunk_dims[:rank])
        out = np.zeros(self.shape, dtype=self.dtype.newbyteorder("="))
        if btree == _UNDEF:
            return out
        for size, mask, offsets, address in _chunk_leaves(self._f._r, btree, rank):
            raw = self._f._r.b[address:address + size]
            for i, (fid, cd) in reversed(list(enumerate(self._filters))):
                if mask & (1 << i):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    es = cd[0] if cd else self.dtype.itemsize
                    a = np.frombuffer(raw, dtype=np.uint8)
                    m = a.size // es
                    raw = a[:m * es].reshape(es, m).T.tobytes() + a[m * es:].tobytes()
                elif fid == 3:
                    raw = raw[:-4]
                else:
                    raise H5Unsupported(f"{self.name}: filter {fid} is not supported")
            chunk = np.frombuffer(raw, dtype=self.dtype, count=int(np.prod(cshape))).reshape(cshape)
            sel_out = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offsets, cshape, self.shape))
            sel_in = tuple(slice(0, s.stop - s.start) for s in sel_out)
            out[sel_out] = chunk[sel_in]
        return out


class Group:
    def __init__(self, f: "File", name: str, btree: int, heap: int):
        self._f, self.name = f, name
        self._links: Dict[str, int] = {}
        r = f._r
        if r.b[heap:heap + 4] != b"HEAP":
            raise H5Unsupported(f"{name}: bad local heap")
        heap_data = r.addr(heap + 8 + 2 * r.L)
        for name_off, header in _group_entries(r, btree):
            end = r.b.index(b"\x00", heap_data + name_off)
            self._links[r.b[heap_data + name_off:end].decode()] = header

    def keys(self) -> List[str]:
        return list(self._links.keys())

    def __contains__(self, name: str) -> bool:
        return name in self._links

    def __iter__(self):
        return iter(self._links)

    def __len__(self) -> int:
        return len(self._links)

    def __getitem__(self, path: str):
        node = self
        for part in [p for p in path.split("/") if p]:
            if not isinstance(node, Group) or part not in node._links:
                raise KeyError(path)
            node = node._f._object(node._links[part], (node.name.rstrip("/") + "/" + part))
        return node


class File(Group):
    """`File(path)` / `File(path, "r")`: the root group of an HDF5 file held in memory."""

    def __init__(self, path: str, mode: str = "r"):
        if mode != "r":
            raise ValueError("h5lite is read-only")
        with open(path, "rb") as fh:
            buf = fh.read()
        if buf[:8] != _SIG:
            raise ValueError(f"{path}: not an HDF5 file")
        r = self._r = _Reader(buf)
        version = buf[8]
        if version not in (0, 1):
            raise H5Unsupported(f"{path}: superblock version {version} (written with libver='latest'?) is not supported")
        r.O, r.L = buf[13], buf[14]
        pos = 24 + (4 if version == 1 else 0)
        base = r.addr(pos)
        if base not in (0, _UNDEF):
            raise H5Unsupported(f"{path}: non-zero base address")
        root = pos + 4 * r.O                      # root group symbol-table entry
        header = r.addr(root + r.O)
        cache_type = r.u(root + 2 * r.O, 4)
        self._cache: Dict[int, object] = {}
        if cache_type == 1:
            scratch = root + 2 * r.O + 8
            Group.__init__(self, self, "/", r.addr(scratch), r.addr(scratch + r.O))
        else:
            obj = self._object(header, "/")
            if not isinstance(obj, Group):
                raise H5Unsupported(f"{path}: root object is not a group")
            Group.__init__(self, self, "/", *obj._addresses)
        self.filename = path

    def _object(self, header: int, name: str):
        hit = self._cache.get(header)
        if hit is not None:
            return hit
        msgs = _header_messages(self._r, header, name)
        stab = [m for m in msgs if m[0] == 0x11]
        if stab:
            off = stab[0][1]
            obj = Group(self, name, self._r.addr(off), self._r.addr(off + self._r.O))
            obj._addresses = (self._r.addr(off), self._r.addr(off + self._r.O))
        else:
            obj = Dataset(self, name, msgs)
        self._cache[header] = obj
        return obj

    def close(self):
        self._r = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


# ------------------------------------------------------------------------------------------------- format pieces
def _header_messages(r: _Reader, header: int, name: str) -> List[Tuple[int, int, int]]:
    """[(message type, data offset, data size)] of a version-1 object header, continuation blocks followed."""
    if r.b[header:header + 4] == b"OHDR":
        raise H5Unsupported(f"{name}: version-2 object header (file written with libver='latest') is not supported")
    if r.b[header] != 1:
        raise H5Unsupported(f"{name}: object header version {r.b[header]}")
    total = r.u(header + 2, 2)
    blocks = [(header + 16, r.u(header + 8, 4))]
    msgs: List[Tuple[int, int, int]] = []
    while blocks and len(msgs) < total:
        pos, length = blocks.pop(0)
        end = pos + length
        while pos + 8 <= end and len(msgs) < total:
            mtype, size = r.u(pos, 2), r.u(pos + 2, 2)
            data = pos + 8
            if mtype == 0x10:
                blocks.append((r.addr(data), r.u(data + r.O, r.L)))
            msgs.append((mtype, data, size))
            pos = data + size
    return msgs


def _dataspace(r: _Reader, off: int) -> Tuple[int, ...]:
    version, rank = r.b[off], r.b[off + 1]
    if version == 1:
        start = off + 8
    elif version == 2:
        if r.b[off + 3] == 2:         # null dataspace
            return (0,)
        start = off + 4
    else:
        raise H5Unsupported(f"dataspace version {version}")
    return tuple(r.u(start + i * r.L, r.L) for i in range(rank))


def _datatype(r: _Reader, off: int) -> np.dtype:
    cls, bits0, size = r.b[off] & 0x0F, r.b[off + 1], r.u(off + 4, 4)
    order = ">" if bits0 & 1 else "<"
    if cls == 0:
        kind = "i" if bits0 & 0x08 else "u"
    elif cls == 1:
        kind = "f"
    else:
        raise H5Unsupported(f"datatype class {cls} (only integers and IEEE floats are supported)")
    if size not in (1, 2, 4, 8) or (kind == "f" and size == 1):
        raise H5Unsupported(f"datatype size {size}")
    return np.dtype(f"{order}{kind}{size}")


def _layout(r: _Reader, off: int):
    version = r.b[off]
    if version != 3:
        raise H5Unsupported(f"data layout message version {version}")
    cls = r.b[off + 1]
    if cls == 0:
        size = r.u(off + 2, 2)
        return "compact", r.b[off + 4:off + 4 + size]
    if cls == 1:
        return "contiguous", (r.addr(off + 2), r.u(off + 2 + r.O, r.L))
    if cls == 2:
        rank = r.b[off + 2]
        btree = r.addr(off + 3)
        dims = [r.u(off + 3 + r.O + 4 * i, 4) for i in range(rank)]
        return "chunked", (btree, dims)
    raise H5Unsupported(f"data layout class {cls}")


def _filters(r: _Reader, off: int) -> List[Tuple[int, List[int]]]:
    version, n = r.b[off], r.b[off + 1]
    pos = off + (8 if version == 1 else 2)
    out = []
    for _ in range(n):
        fid = r.u(pos, 2)
        pos += 2
        name_len = 0
        if version == 1 or fid >= 256:
            name_len = r.u(pos, 2)
            pos += 2
        pos += 2                                  # flags
        ncd = r.u(pos, 2)
        pos += 2
        pos += (name_len + 7) // 8 * 8 if version == 1 else name_len
        cd = [r.u(pos + 4 * i, 4) for i in range(ncd)]
        pos += 4 * ncd
        if version == 1 and ncd % 2:
            pos += 4
        out.append((fid, cd))
    return out


def _group_entries(r: _Reader, node: int):
    """(heap offset of the link name, object header address) of every link under a version-1 group B-tree node."""
    sig = r.b[node:node + 4]
    if sig == b"SNOD":
        count = r.u(node + 6, 2)
        entry = 2 * r.O + 24
        for i in range(count):
            e = node + 8 + i * entry
            yield r.u(e, r.O), r.addr(e + r.O)
        return
    if sig != b"TREE" or r.b[node + 4] != 0:
        raise H5Unsupported("group B-tree node expected")
    used = r.u(node + 6, 2)
    pos = node + 8 + 2 * r.O
    for i in range(used):
        child = r.addr(pos + r.L + i * (r.L + r.O))
        yield from _group_entries(r, child)


def _chunk_leaves(r: _Reader, node: int, rank: int):
    """(stored size, filter mask, element offsets, address) of every chunk under a version-1 chunk B-tree node."""
    if r.b[node:node + 4] != b"TREE" or r.b[node + 4] != 1:
        raise H5Unsupported("chunk B-tree node expected")
    level, used = r.b[node + 5], r.u(node + 6, 2)
    key = 8 + 8 * (rank + 1)
    pos = node + 8 + 2 * r.O
    for i in range(used):
        k = pos + i * (key + r.O)
        child = r.addr(k + key)
        if level > 0:
            yield from _chunk_leaves(r, child, rank)
        else:
            yield (r.u(k, 4), r.u(k + 4, 4), tuple(r.u(k + 8 + 8 * d, 8) for d in range(rank)), child)
