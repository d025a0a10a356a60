"""GPU: bit-exact parity of the device graph set-up (libmgn_b200.so through
graph_physics_b200.graph) with the oracle and the committed reference fixtures."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import graph_oracle as go

pytestmark = pytest.mark.gpu


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _dev(a, dtype=torch.int64):
    return torch.as_tensor(np.asarray(a), dtype=dtype).cuda()


def test_cylinder_edges_khop_csr(golden_dir):
    from graph_physics_b200 import graph as G
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n = z["points"].shape[0]
    face = G.cells_to_faces(_dev(z["cells"]))
    assert np.array_equal(face.cpu().numpy(), z["face"])
    ei = G.face_to_edge(face, n)
    assert ei.dtype == torch.int64 and tuple(ei.shape) == (2, 11070)
    assert np.array_equal(ei.cpu().numpy(), z["edge_index"])
    k2 = G.compute_k_hop_edge_index(ei, 2, n)
    assert tuple(k2.shape) == (2, 32638) and np.array_equal(k2.cpu().numpy(), z["khop2"])
    with open(os.path.join(golden_dir, "meta.json")) as f:
        m = json.load(f)["cylinder"]
    k3 = G.compute_k_hop_edge_index(ei, 3, n)
    assert k3.shape[1] == m["E_khop3"] and _sha(k3.cpu().numpy()) == m["khop3_sha256"]
    csr = G.build_csr(ei, n)
    rowptr, perm, send, recv = go.receiver_csr(z["edge_index"].astype(np.int64), n)
    assert np.array_equal(csr.rowptr.cpu().numpy(), rowptr)
    assert np.array_equal(csr.perm.cpu().numpy(), perm)
    assert np.array_equal(csr.senders.cpu().numpy(), send)
    assert np.array_equal(csr.receivers.cpu().numpy(), recv)
    ea = G.edge_features(_dev(z["points"], torch.float32), ei)
    ref = go.edge_features(z["points"], z["edge_index"].astype(np.int64))
    assert tuple(ea.shape) == (11070, 4)
    np.testing.assert_allclose(ea.cpu().numpy(), ref, rtol=1e-6, atol=1e-7)
    und = G.to_undirected(ei[:, ::2].contiguous(), n)
    assert np.array_equal(und.cpu().numpy(), go.to_undirected(z["edge_index"][:, ::2].astype(np.int64), n))


def test_aneurysm_tets(golden_dir):
    from graph_physics_b200 import graph as G
    z = np.load(os.path.join(golden_dir, "aneurysm_mesh.npz"))
    with open(os.path.join(golden_dir, "meta.json")) as f:
        m = json.load(f)["aneurysm"]
    n = z["points"].shape[0]
    face = G.cells_to_faces(_dev(z["cells"]))
    assert _sha(face.cpu().numpy()) == m["face_sha256"]
    ei = G.face_to_edge(face, n)
    assert ei.shape[1] == 291144 and _sha(ei.cpu().numpy()) == m["edge_index_sha256"]
    k2 = G.compute_k_hop_edge_index(ei, 2, n)
    assert k2.shape[1] == 1395256 and _sha(k2.cpu().numpy()) == m["khop2_sha256"]
    csr = G.build_csr(k2, n)
    rowptr, perm, send, recv = go.receiver_csr(k2.cpu().numpy(), n)
    assert np.array_equal(csr.rowptr.cpu().numpy(), rowptr) and np.array_equal(csr.perm.cpu().numpy(), perm)
    assert np.array_equal(csr.senders.cpu().numpy(), send)


def test_meshmask_known_answers_and_reference_run(golden_dir):
    from graph_physics_b200 import graph as G
    # reference tests/graphphysics/utils/test_meshmask.py:59-120
    ei = _dev([[0, 1, 1, 2, 3, 4], [1, 0, 2, 1, 4, 3]])
    ea = torch.arange(48, dtype=torch.float32, device="cuda").view(6, 8)
    f_ei, f_ea, mask = G.filter_edges(ei, _dev([0, 1, 2]), ea)
    assert f_ei.cpu().tolist() == [[0, 1, 1, 2], [1, 0, 2, 1]]
    assert mask.cpu().tolist() == [True, True, True, True, False, False] and mask.dtype == torch.bool
    assert torch.equal(f_ea, ea[:4])
    f_ei, f_ea, mask = G.filter_edges(ei, _dev([0, 3]), ea)
    assert tuple(f_ei.shape) == (2, 0) and tuple(f_ea.shape) == (0, 8) and not mask.any()
    f_ei, f_ea, mask = G.filter_edges(_dev([[0, 1], [1, 0]]), _dev([0, 1]), None)
    assert f_ei.cpu().tolist() == [[0, 1], [1, 0]] and f_ea is None and mask.all()
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    mm = np.load(os.path.join(golden_dir, "meshmask_cylinder.npz"))
    ids = torch.arange(11070, dtype=torch.int32, device="cuda").view(-1, 1)
    f_ei, f_ea, mask = G.filter_edges(_dev(z["edge_index"]), _dev(mm["selected"]), ids)
    assert np.array_equal(f_ei.cpu().numpy(), mm["edge_index"])
    assert np.array_equal(mask.cpu().numpy(), mm["mask"])
    assert np.array_equal(f_ea.cpu().numpy().ravel(), mm["kept_ids"])


def test_empty_and_tile_batch():
    from graph_physics_b200 import graph as G
    e = G.face_to_edge(torch.empty((3, 0), dtype=torch.int64, device="cuda"), 5)
    assert tuple(e.shape) == (2, 0)
    csr = G.build_csr(e, 5)
    assert csr.rowptr.cpu().tolist() == [0] * 6
    ei = _dev([[0, 1, 1, 2], [1, 0, 2, 1]])
    b = G.tile_batch(ei, 3, 3)
    assert b.cpu().tolist() == [[0, 1, 1, 2, 3, 4, 4, 5, 6, 7, 7, 8], [1, 0, 2, 1, 4, 3, 5, 4, 7, 6, 8, 7]]
