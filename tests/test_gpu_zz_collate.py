"""GPU: the CSR of a collated batch assembled from the samples' device-built CSRs (`batch.concat_csr`) is bit-identical
to `mgn_build_csr` run on the batch's own `edge_index`, and the processor picks the registered CSR up.

(The logic is device independent and covered by tests/test_cpu_batch.py against the oracle; this file checks it on the
device-built structures.)"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_concat_csr_equals_device_build(golden_dir):
    from graph_physics_b200 import batch as B, graph as G, processor as proc
    from graph_physics_b200.models import Graph
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    dev = "cuda"
    ei_mesh = torch.as_tensor(z["edge_index"].astype(np.int64), device=dev)
    n_mesh = int(z["points"].shape[0])
    g = torch.Generator(device="cpu").manual_seed(9)
    samples = []
    for n, ei in ((n_mesh, ei_mesh), (300, torch.randint(0, 300, (2, 1000), generator=g).to(dev)), (n_mesh, ei_mesh)):
        samples.append(Graph(x=torch.randn(n, 128, generator=g).to(dev),
                             edge_attr=torch.randn(ei.shape[1], 128, generator=g).to(dev), edge_index=ei))
    csrs = [proc.csr_for(s.edge_index, s.x.shape[0]) for s in samples]
    b = B.collate(samples, csrs=csrs)
    want = G.build_csr(b.edge_index, b.num_nodes)
    for name in ("rowptr", "perm", "senders", "receivers"):
        assert torch.equal(getattr(b.csr, name), getattr(want, name)), name
    assert b.csr.identity == want.identity
    assert proc.csr_for(b.edge_index, b.num_nodes) is b.csr


def test_resident_trajectory_dataset_serves_the_reference_frames(golden_dir, tmp_path):
    """`TrajectoryDataset` (device-resident trajectories) against the host functions that tests/test_cpu_dataset.py pins
    to the reference: sample i is frame (i mod T-1) with target frame + 1; one edge_index object and one CSR per
    trajectory; a batch of samples collates without a sort."""
    import json
    from graph_physics_b200 import batch as B, dataset as D, graph as G
    z = np.load(os.path.join(golden_dir, "ref_dataset.npz"))
    meta = json.loads(bytes(z["cyl2d.meta"]).decode())
    group = {k[len("cyl2d.raw."):]: z[k] for k in z.files if k.startswith("cyl2d.raw.")}
    path = os.path.join(tmp_path, "train.npz")
    D.save_npz_container(path, {"t0": group, "t1": group})
    meta_path = os.path.join(tmp_path, "meta.json")
    with open(meta_path, "w") as fp:
        json.dump(meta, fp)
    ds = D.TrajectoryDataset(path, meta_path, device="cuda")
    traj = D.traj_as_meshes(group, meta)
    T = meta["trajectory_length"]
    assert len(ds) == 2 * (T - 1)
    samples = [ds[i] for i in (0, 3, T - 1 + 2)]
    for s, frame in zip(samples, (0, 3, 2)):
        want = D.frame_as_graph(traj, frame, meta, frame_target=frame + 1)
        assert torch.equal(s.x.cpu(), want.x) and torch.equal(s.y.cpu(), want.y) and torch.equal(s.pos.cpu(), want.pos)
        ei = G.face_to_edge(want.face.cuda().contiguous(), want.x.shape[0])
        assert torch.equal(s.edge_index, ei)
        assert torch.equal(s.edge_attr, G.edge_features(want.pos.cuda(), ei))
        assert s.csr.num_edges == ei.shape[1] and not hasattr(s, "previous_data")
    assert samples[0].edge_index is samples[1].edge_index and samples[0].csr is samples[1].csr     # same trajectory
    assert samples[0].traj_index == 0 and samples[2].traj_index == 1
    b = B.collate(samples, csrs=[s.csr for s in samples])
    want = G.build_csr(b.edge_index, b.num_nodes)
    assert torch.equal(b.csr.perm, want.perm) and torch.equal(b.csr.rowptr, want.rowptr)
    ds.close()


def test_preprocessing_pipeline_on_device(golden_dir):
    """`preprocessing.build_preprocessing` on a raw sample (x, y, pos, face): the composed graph equals the graph.py
    functions called by hand, with and without world edges."""
    from graph_physics_b200 import graph as G, preprocessing as P
    from graph_physics_b200.models import Graph
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    dev = "cuda"
    pos = torch.as_tensor(z["points"][:, :2]).to(dev)
    face = torch.as_tensor(z["face"].astype(np.int64)).to(dev)
    n = pos.shape[0]
    g = torch.Generator().manual_seed(2)
    x = torch.randn(n, 5, generator=g)
    x[:, 4] = torch.randint(0, 2, (n,), generator=g).float()
    x[:, :2] = torch.as_tensor(z["points"][:, :2]) + 0.01 * torch.randn(n, 2, generator=g)
    x, y = x.to(dev), torch.randn(n, 2, generator=g).to(dev)
    out = P.build_preprocessing()(Graph(x=x, y=y, pos=pos, face=face))
    ei = G.face_to_edge(face, n)
    assert torch.equal(out.edge_index, ei) and torch.equal(out.edge_attr, G.edge_features(pos, ei))
    nz = {"noise_index_start": 0, "noise_index_end": 2, "noise_scale": 0.02, "node_type_index": 4}
    noisy = P.build_preprocessing(noise_parameters=nz)(Graph(x=x, y=y, pos=pos, face=face))
    moved = (noisy.x != x).any(dim=1)
    assert torch.equal(moved, x[:, 4] == 0) and torch.equal(noisy.edge_index, ei)
    ei_w = G.add_world_edges(ei, x[:, :2].contiguous(), x[:, 4], n, radius=0.05)
    w = P.add_world_edges(Graph(x=x, edge_index=ei, edge_attr=out.edge_attr), 0, 2, 4, radius=0.05)
    assert torch.equal(w.edge_index, ei_w) and w.edge_attr is None
    w = P.add_world_pos_features(P.edge_features(Graph(x=x, pos=pos, edge_index=ei_w, edge_attr=None)), 0, 2)
    assert w.edge_attr.shape == (ei_w.shape[1], 6)


def test_k_hop_is_applied_after_the_preprocessing(golden_dir, tmp_path):
    """`BaseDataset.__getitem__` runs the preprocessing (which may add world edges) and THEN `_apply_k_hop`
    (graphphysics/dataset/h5_dataset.py:100-101, dataset.py:139-175): with world-edge preprocessing and khop = 2 the
    edge set is the 2-hop expansion of mesh + world edges (oracle: numpy restatement), cached per trajectory; without a
    preprocessing the expansion of the mesh edges comes from the resident topology (same as before)."""
    import json
    from graph_physics_b200 import dataset as D, preprocessing as P
    from oracle import graph_oracle as go
    z = np.load(os.path.join(golden_dir, "ref_dataset.npz"))
    meta = json.loads(bytes(z["cyl2d.meta"]).decode())
    group = {k[len("cyl2d.raw."):]: z[k] for k in z.files if k.startswith("cyl2d.raw.")}
    path = os.path.join(tmp_path, "train.npz")
    D.save_npz_container(path, {"t0": group})
    meta_path = os.path.join(tmp_path, "meta.json")
    with open(meta_path, "w") as fp:
        json.dump(meta, fp)
    plain = D.TrajectoryDataset(path, meta_path, device="cuda", khop=2)
    s0 = plain[0]
    n = s0.x.shape[0]
    ei1 = go.face_to_edge(s0.face.cpu().numpy(), n)
    assert np.array_equal(s0.edge_index.cpu().numpy(), go.k_hop_edge_index(ei1, 2, n))
    plain.close()
    # world edges between "obstacle" and "normal" nodes.  Column layout after the transforms (the reference's
    # DeformingPlate convention, preprocessing.py:47-89): [world_pos (3) | obstacle displacement (3) | node type | ...];
    # node types are made up from the position, the world position is the mesh position with z = 0
    def add_cols(g):
        n_ = g.pos.shape[0]
        nt_ = (g.pos[:, 0] > g.pos[:, 0].median()).to(g.x.dtype).unsqueeze(1)          # 1 = OBSTACLE, 0 = NORMAL
        wpos = torch.cat([g.pos.to(g.x.dtype), torch.zeros(n_, 1, dtype=g.x.dtype, device=g.x.device)], dim=1)
        g.x = torch.cat([wpos, nt_, g.x], dim=1)
        g.y = torch.cat([wpos + 0.001, g.y], dim=1)
        return g
    wp = {"world_pos_index_start": 0, "world_pos_index_end": 3, "node_type_index": 6, "radius": 0.06}
    pre = P.build_preprocessing(world_pos_parameters=wp, extra_node_features=add_cols)
    ds = D.TrajectoryDataset(path, meta_path, device="cuda", khop=2, preprocessing=pre, world_pos_index_start=0,
                             world_pos_index_end=3)
    s = ds[0]
    wpos = s.x[:, 0:3].cpu().numpy()
    nt = s.x[:, 6].cpu().numpy()
    ei_world = go.world_edges(ei1, wpos.astype(np.float64), nt, n, radius=0.06)
    assert ei_world.shape[1] > ei1.shape[1], "the fixture must produce world edges for this test to mean anything"
    want = go.k_hop_edge_index(ei_world, 2, n)
    assert np.array_equal(s.edge_index.cpu().numpy(), want)
    assert not np.array_equal(want, go.k_hop_edge_index(ei1, 2, n))
    assert s.edge_attr.shape == (want.shape[1], 7)                 # Cartesian + Distance (3) + world-position features (4)
    assert ds[1].edge_index is s.edge_index                        # cached per trajectory, as the reference does
    ds.close()
