"""CPU: the C-ABI library builds, loads and exports every symbol include/mgn_b200.h
declares; the product path refuses to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from graph_physics_b200 import build
    path = build.build()
    lib = ctypes.CDLL(path)
    with open(os.path.join(ROOT, "include", "mgn_b200.h")) as f:
        hdr = f.read()
    names = sorted(set(re.findall(r"\b(mgn_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"libmgn_b200.so does not export {n}"
    assert lib.mgn_version() >= 100


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU refusal")
def test_no_cpu_fallback():
    from graph_physics_b200 import _lib, graph
    with pytest.raises(_lib.MgnError):
        graph.face_to_edge(torch.zeros((3, 4), dtype=torch.int64), 4)
