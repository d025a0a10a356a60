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


def test_product_modules_never_import_the_oracle():
    """oracle/ is test infrastructure: no module of the product package may import, load or execute it.
    Checked on the source text of every product file and on sys.modules after importing the whole package."""
    import importlib
    import pkgutil
    import subprocess
    import sys
    pkg_dir = os.path.join(ROOT, "graph_physics_b200")
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|import_module\(\s*[\"']oracle|[\"']oracle[/.]", re.M)
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, f)) as fh:
                    assert not pat.search(fh.read()), f"{f} refers to oracle/"
    code = ("import sys, pkgutil, importlib; sys.path.insert(0, %r); import graph_physics_b200 as g\n"
            "for m in pkgutil.walk_packages(g.__path__, 'graph_physics_b200.'):\n"
            "    if not m.name.rsplit('.', 1)[-1].startswith('lib'):\n"
            "        importlib.import_module(m.name)\n"
            "bad = [m for m in sys.modules if m == 'oracle' or m.startswith('oracle.')]\n"
            "assert not bad, bad\n") % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
