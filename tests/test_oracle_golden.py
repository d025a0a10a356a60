"""CPU: pins the oracle (oracle/) to the fixtures that tests/golden/make_golden.py
produced by running the reference's own source, and to the reference's own golden
counts / known-answer tensors."""
import hashlib
import json
import os

import numpy as np
import torch

from oracle import graph_oracle as go
from oracle import mgn_oracle as mo
from oracle.synth import processor_inputs


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _meta(golden_dir):
    with open(os.path.join(golden_dir, "meta.json")) as f:
        return json.load(f)


def test_cylinder_edge_counts_match_reference_goldens(golden_dir):
    # reference tests/graphphysics/dataset/test_xdmfdataset.py:171-175, 226-231
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n = z["points"].shape[0]
    face = go.cells_to_faces(z["cells"])
    assert np.array_equal(face, z["face"])
    ei = go.face_to_edge(face, n)
    assert ei.shape == (2, 11070)
    assert np.array_equal(ei, z["edge_index"].astype(np.int64))
    k2 = go.k_hop_edge_index(ei, 2, n)
    assert k2.shape == (2, 32638)
    assert np.array_equal(k2, z["khop2"].astype(np.int64))
    k3 = go.k_hop_edge_index(ei, 3, n)
    m = _meta(golden_dir)["cylinder"]
    assert k3.shape[1] == m["E_khop3"] and _sha(k3) == m["khop3_sha256"]
    ea = go.edge_features(z["points"], ei)
    assert ea.shape == (11070, 4)   # test_xdmfdataset.py:175


def test_aneurysm_tet_mesh_hashes(golden_dir):
    z = np.load(os.path.join(golden_dir, "aneurysm_mesh.npz"))
    m = _meta(golden_dir)["aneurysm"]
    n = z["points"].shape[0]
    face = go.cells_to_faces(z["cells"])
    assert _sha(face.astype(np.int64)) == m["face_sha256"]
    ei = go.face_to_edge(face, n)
    assert ei.shape[1] == m["E"] == 291144
    assert _sha(ei) == m["edge_index_sha256"]
    k2 = go.k_hop_edge_index(ei, 2, n)
    assert k2.shape[1] == m["E_khop2"] == 1395256
    assert _sha(k2) == m["khop2_sha256"]


def test_meshmask_known_answers():
    # reference tests/graphphysics/utils/test_meshmask.py:21-23, 59-120
    ei = np.array([[0, 1, 1, 2, 3, 4], [1, 0, 2, 1, 4, 3]])
    ea = np.arange(48, dtype=np.float32).reshape(6, 8)
    f_ei, f_ea, mask = go.filter_edges(ei, np.array([0, 1, 2]), ea)
    assert np.array_equal(f_ei, [[0, 1, 1, 2], [1, 0, 2, 1]])
    assert np.array_equal(mask, [True, True, True, True, False, False])
    assert np.array_equal(f_ea, ea[:4])
    f_ei, f_ea, mask = go.filter_edges(ei, np.array([0, 3]), ea)
    assert f_ei.shape == (2, 0) and f_ea.shape == (0, 8) and not mask.any()
    f_ei, f_ea, mask = go.filter_edges(np.array([[0, 1], [1, 0]]), np.array([0, 1]), None)
    assert np.array_equal(f_ei, [[0, 1], [1, 0]]) and f_ea is None and mask.all()
    g, keep = go.build_masked_graph(np.arange(10.).reshape(5, 2), None, ei, ea, np.array([0, 1, 2]))
    assert g["x"].shape == (3, 2) and np.array_equal(g["edge_index"], [[0, 1, 1, 2], [1, 0, 2, 1]])
    feats, eattr = go.reconstruct_graph((5, 2), np.ones((3, 2), np.float32), np.array([0, 1, 2]),
                                        np.array([[0.5, -0.5]], np.float32), ea, np.zeros((4, 8), np.float32),
                                        keep, np.full((1, 8), 2.0, np.float32))
    assert np.array_equal(feats[3:], [[0.5, -0.5]] * 2) and np.array_equal(feats[:3], np.ones((3, 2)))
    assert np.array_equal(eattr[:4], np.zeros((4, 8))) and np.array_equal(eattr[4:], ea[4:] + 2.0)


def test_meshmask_matches_reference_run(golden_dir):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    mm = np.load(os.path.join(golden_dir, "meshmask_cylinder.npz"))
    ei = z["edge_index"].astype(np.int64)
    ids = np.arange(ei.shape[1], dtype=np.int32)[:, None]
    f_ei, f_ea, mask = go.filter_edges(ei, mm["selected"], ids)
    assert np.array_equal(f_ei, mm["edge_index"]) and np.array_equal(mask, mm["mask"])
    assert np.array_equal(f_ea.ravel(), mm["kept_ids"])


def test_random_edges_count_matches_reference_golden(golden_dir):
    # test_xdmfdataset.py:207-212: 11,070 -> 16,604 at p=0.5
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    ei = z["edge_index"].astype(np.int64)
    out, add = go.add_random_edges(ei, 0.5, 1923, np.random.default_rng(0))
    assert out.shape[1] == 16604
    assert np.array_equal(out[:, :11070], ei)
    key = set((add[0] * 1923 + add[1]).tolist())
    assert all((b * 1923 + a) in key for a, b in add.T.tolist())


def test_receiver_csr_properties(golden_dir):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    ei = z["edge_index"].astype(np.int64)
    rowptr, perm, send, recv = go.receiver_csr(ei, 1923)
    assert rowptr[0] == 0 and rowptr[-1] == 11070 and (np.diff(recv) >= 0).all()
    assert np.array_equal(ei[1][perm], recv) and np.array_equal(ei[0][perm], send)
    for nd in (0, 7, 1922):
        assert (recv[rowptr[nd]:rowptr[nd + 1]] == nd).all()


def test_mlp_and_rmsnorm_match_reference_run(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_mlp_rmsnorm.npz"))
    p = {k: torch.tensor(z[k]) for k in z.files if k.startswith("m.")}
    x = torch.tensor(z["x"])
    torch.testing.assert_close(mo.mlp(x, p, "m"), torch.tensor(z["mlp_out"]), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(mo.rms_norm(x, torch.ones(24)), torch.tensor(z["rms_out"]), rtol=1e-6, atol=1e-7)


def test_toy_ring_matches_reference_run(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_toy_ring.npz"))
    p = mo.init_params(3, hidden=16, seed=3)
    x, e = torch.tensor(z["x0"]), torch.tensor(z["e0"])
    ei = torch.tensor(z["edge_index"])
    for k in range(3):
        x, e = mo.graph_net_block(x, ei, e, p, f"processor_list.{k}")
        torch.testing.assert_close(x, torch.tensor(z[f"x{k+1}"]), rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(e, torch.tensor(z[f"e{k+1}"]), rtol=1e-5, atol=1e-6)


def test_cylinder_h128_s15_forward_backward_matches_reference_run(golden_dir):
    torch.set_num_threads(os.cpu_count())
    z = np.load(os.path.join(golden_dir, "ref_cylinder_h128_s15.npz"))
    meta = _meta(golden_dir)["cylinder_h128_s15"]
    ei = torch.tensor(np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))["edge_index"].astype(np.int64))
    p = mo.init_params(15, hidden=128, seed=0)
    for v in p.values():
        v.requires_grad_(True)
    x0, e0, gx = processor_inputs(meta["N"], meta["E"], 128, seed=0)
    x0.requires_grad_(True)
    e0.requires_grad_(True)
    x, e, xs, es = mo.processor(x0, ei, e0, p, 15, return_all=True)
    assert mo.rel_l2(x, torch.tensor(z["x15"])) < 1e-6
    assert mo.rel_l2(xs[0], torch.tensor(z["x1"])) < 1e-6
    np.testing.assert_allclose([float(t.double().norm()) for t in xs], z["x_norms"], rtol=1e-6)
    np.testing.assert_allclose([float(t.double().norm()) for t in es], z["e_norms"], rtol=1e-6)
    (x * gx).sum().backward()
    assert mo.rel_l2(x0.grad, torch.tensor(z["gx0"])) < 1e-5
    for name, gn in meta["grad_norms"].items():
        assert abs(float(p[name].grad.double().norm()) - gn) <= 1e-4 * gn + 1e-12, name
    for key in z.files:
        if key.startswith("g:"):
            assert mo.rel_l2(p[key[2:]].grad, torch.tensor(z[key])) < 1e-4, key


def test_epd_small_matches_reference_run(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_epd_small.npz"))
    ei = torch.tensor(np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))["edge_index"].astype(np.int64))
    p = mo.init_params(3, hidden=32, node_in=11, edge_in=4, out_size=2, seed=1)
    y = mo.encode_process_decode(torch.tensor(z["graph_x"]), ei, torch.tensor(z["edge_attr"]), p, 3)
    assert mo.rel_l2(y, torch.tensor(z["out"])) < 1e-5
