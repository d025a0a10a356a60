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


def _world_case(n, dim, radius, seed, clustered=False):
    rng = np.random.default_rng(seed)
    pos = rng.random((n, dim)).astype(np.float32)
    if clustered:      # an obstacle blob touching a plate, like DeformingPlate
        pos[: n // 4] = (0.5 + 0.08 * rng.standard_normal((n // 4, dim))).astype(np.float32)
    typ = rng.choice(np.array([0, 1, 3]), size=n, p=[0.6, 0.25, 0.15])        # NORMAL, OBSTACLE, HANDLE
    if clustered:
        typ[: n // 4] = 1
    mesh = rng.integers(0, n, size=(2, 3 * n))
    mesh = mesh[:, mesh[0] != mesh[1]]
    return pos, typ, go.to_undirected(mesh.astype(np.int64), n)


@pytest.mark.parametrize("n,dim,radius,clustered", [(3000, 3, 0.05, False), (2500, 2, 0.02, False),
                                                   (4000, 3, 0.03, True), (50, 3, 2.0, False)])
def test_world_edges_equal_ckdtree_path(n, dim, radius, clustered):
    """S4 add_world_edges (preprocessing.py:92-140): the device cell list finds exactly the pairs of the
    cKDTree.query_pairs + node-type-mask path (oracle restatement, scipy), bit-exact after to_undirected."""
    from graph_physics_b200 import graph as G
    pos, typ, mesh = _world_case(n, dim, radius, seed=n + dim, clustered=clustered)
    want = go.world_edges(mesh, pos, typ, n, radius=radius)
    got = G.add_world_edges(_dev(mesh), torch.as_tensor(pos).cuda(), torch.as_tensor(typ).cuda(), n, radius=radius)
    assert want.shape[1] > mesh.shape[1]                     # the case does add world edges
    assert np.array_equal(got.cpu().numpy(), want)
    # the reference reads the node type from a float feature column
    got_f = G.add_world_edges(_dev(mesh), torch.as_tensor(pos).cuda(), torch.as_tensor(typ).float().cuda(), n,
                              radius=radius)
    assert torch.equal(got_f, got)
    # pairs: obstacle first, normal second, every one within the radius
    pr = G.world_pairs(torch.as_tensor(pos).cuda(), torch.as_tensor(typ).cuda(), radius).cpu().numpy()
    assert (typ[pr[0]] == 1).all() and (typ[pr[1]] == 0).all()
    d = np.linalg.norm(pos[pr[0]].astype(np.float64) - pos[pr[1]].astype(np.float64), axis=1)
    assert (d <= radius).all() and len({(a, b) for a, b in pr.T}) == pr.shape[1]


def test_world_edges_per_sample_and_degenerate():
    from graph_physics_b200 import graph as G
    # two samples overlaid in space: pairs must not cross samples
    pos1, typ1, mesh1 = _world_case(800, 3, 0.08, seed=1)
    pos2, typ2, mesh2 = _world_case(700, 3, 0.08, seed=2)
    w1 = go.world_edges(mesh1, pos1, typ1, 800, radius=0.08)
    w2 = go.world_edges(mesh2, pos2, typ2, 700, radius=0.08)
    want = np.concatenate([w1, w2 + 800], axis=1)
    pos = torch.as_tensor(np.concatenate([pos1, pos2])).cuda()
    typ = torch.as_tensor(np.concatenate([typ1, typ2])).cuda()
    mesh = _dev(np.concatenate([mesh1, mesh2 + 800], axis=1))
    sample = torch.cat([torch.zeros(800), torch.ones(700)]).int().cuda()
    got = G.add_world_edges(mesh, pos, typ, 1500, radius=0.08, sample=sample)
    assert np.array_equal(got.cpu().numpy(), want)
    # no obstacle nodes: the mesh edges come back (canonicalised)
    none = G.add_world_edges(_dev(mesh1), torch.as_tensor(pos1).cuda(), torch.zeros(800).cuda(), 800, radius=0.08)
    assert np.array_equal(none.cpu().numpy(), mesh1)
    # all points identical (one cell), zero nodes
    same = torch.zeros(40, 3).cuda()
    t = torch.as_tensor([1] * 5 + [0] * 35).cuda()
    assert G.world_pairs(same, t, 0.03).shape[1] == 5 * 35
    assert G.world_pairs(torch.zeros(0, 3).cuda(), torch.zeros(0).cuda(), 0.03).shape == (2, 0)
    with pytest.raises(ValueError):
        G.world_pairs(same, t, 0.0)


def test_random_edges_count_symmetry_disjointness(golden_dir):
    """S5 _add_random_edges (dataset.py:104-137): the reference's own fixture count (11,070 -> 16,604 at p = 0.5,
    test_xdmfdataset.py:207-212), both directions present, nothing duplicated, originals untouched and first."""
    from graph_physics_b200 import graph as G
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n = int(z["points"].shape[0])
    ei = _dev(z["edge_index"])
    gen = torch.Generator(device="cuda").manual_seed(0)
    out, added = G.add_random_edges(ei, 0.5, n, generator=gen)
    assert out.shape[1] == 16604 and added.shape[1] == 16604 - 11070
    assert torch.equal(out[:, :11070], ei)
    k = (out[0] * n + out[1]).cpu().numpy()
    assert len(np.unique(k)) == k.size                                  # disjoint from the mesh and from each other
    assert bool((added[0] != added[1]).all())
    kr = (added[1] * n + added[0]).cpu().numpy()
    assert np.array_equal(np.sort(kr), np.sort((added[0] * n + added[1]).cpu().numpy()))   # symmetric
    # the oracle restatement gives the same count with its own RNG
    o_out, _ = go.add_random_edges(z["edge_index"], 0.5, n, np.random.default_rng(0))
    assert o_out.shape[1] == out.shape[1]
    out2, _ = G.add_random_edges(ei, 0.5, n, generator=torch.Generator(device="cuda").manual_seed(1))
    assert not torch.equal(out2, out)
    same, none = G.add_random_edges(ei, 0.0, n)
    assert torch.equal(same, ei) and none.shape[1] == 0
    # a nearly complete graph: only a few absent pairs remain
    m = 12
    full = torch.combinations(torch.arange(m), 2).t().cuda()
    part_ = torch.cat([full[:, :-6], full[:, :-6].flip(0)], dim=1)
    o3, a3 = G.add_random_edges(part_, 0.1, m)
    assert a3.shape[1] == 2 * (int(round(part_.shape[1] * 0.1)) // 2)
    with pytest.raises(ValueError):
        G.add_random_edges(part_, 0.5, m)
