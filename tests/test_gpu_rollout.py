"""GPU: the captured autoregressive rollout (graph_physics_b200/rollout.py) against a literal restatement of the
reference's `_make_prediction` loop (graphphysics/training/lightning_module.py:168-202) run eagerly on the same model."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _literal_loop(predict_fn, graph, frames, o0, o1, nt, prev):
    from graph_physics_b200.rollout import boundary_mask
    last, last_prev, outs = None, None, []
    for fr in frames:
        x_t, y_t = fr[0], fr[1]
        g = graph.clone()
        g.x = x_t.clone()
        if len(fr) > 2 and fr[2] is not None:      # per-frame edges, as the reference's dataset serves them
            g.edge_attr = fr[2]
        if len(fr) > 3 and fr[3] is not None:
            g.edge_index = fr[3]
        if last is not None:
            g.x[:, o0:o1] = last.detach()
            if prev is not None:
                g.x[:, prev[0]:prev[1]] = last_prev.detach()
        mask = boundary_mask(g.x, nt)
        cur = g.x[:, o0:o1]
        with torch.no_grad():
            pred = predict_fn(g)
        pred[mask] = y_t[mask]
        last = pred
        if prev is not None:
            last_prev = pred - cur
        outs.append(pred.clone())
    return torch.stack(outs)


@pytest.mark.parametrize("use_prev", [False, True])
def test_rollout_matches_the_reference_loop(golden_dir, use_prev):
    from graph_physics_b200.models import EncodeProcessDecode, Graph
    from graph_physics_b200.rollout import Rollout
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    ei = torch.as_tensor(z["edge_index"].astype(np.int64)).cuda()
    n, e = z["points"].shape[0], ei.shape[1]
    torch.manual_seed(7)
    F, d, T = 5, 2, 8
    o0, o1, nt = 0, 2, 4
    prev = (2, 4) if use_prev else None
    model = EncodeProcessDecode(4, F, 3, d).cuda().eval()

    def predict(g):       # simulator-like: previous value + network output
        return g.x[:, o0:o1] + 0.1 * model(g)

    node_type = torch.tensor([0, 4, 5, 6])[torch.randint(0, 4, (n,))].float().cuda()
    frames = []
    for _ in range(T):
        x_t = torch.randn(n, F, device="cuda")
        x_t[:, nt] = node_type
        frames.append((x_t, torch.randn(n, d, device="cuda")))
    graph = Graph(x=torch.zeros(n, F, device="cuda"), edge_attr=torch.randn(e, 3, device="cuda"), edge_index=ei)

    want = _literal_loop(predict, graph, frames, o0, o1, nt, prev)
    got_eager = Rollout(predict, graph.clone(), (o0, o1), nt, prev_slice=prev, capture=False).run(frames)
    got_graph = Rollout(predict, graph.clone(), (o0, o1), nt, prev_slice=prev, capture=True).run(frames)
    torch.cuda.synchronize()
    scale = float(want.abs().max())
    # same kernels, same order of operations: only the red.add order of the aggregation differs between runs
    assert float((got_eager - want).abs().max()) <= 1e-3 * scale
    assert float((got_graph - want).abs().max()) <= 1e-3 * scale
    # boundary nodes carry the ground truth at every step
    from graph_physics_b200.rollout import boundary_mask
    m = boundary_mask(frames[0][0], nt)
    assert m.any() and (~m).any()
    for t in range(T):
        assert torch.equal(got_graph[t][m], frames[t][1][m])
    # a second trajectory through the same captured graph
    again = Rollout(predict, graph.clone(), (o0, o1), nt, prev_slice=prev, capture=True)
    a1 = again.run(frames)
    a2 = again.run(frames)
    assert float((a1 - a2).abs().max()) <= 1e-3 * scale


def test_rollout_with_moving_mesh_and_changing_topology(golden_dir):
    """DeformingPlate-style frames: edge features recomputed from moving positions every frame (same topology: copied
    into the static buffer, the captured graph keeps replaying) and, on some frames, a different edge set (world edges
    appear: that step runs eagerly on the frame's own edge_index / edge_attr) -- against the literal reference loop,
    which takes each frame's edges from the dataset (preprocessing.py:401-427, lightning_module.py:168-232)."""
    from graph_physics_b200 import graph as G
    from graph_physics_b200.models import EncodeProcessDecode, Graph
    from graph_physics_b200.rollout import Rollout
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    ei = torch.as_tensor(z["edge_index"].astype(np.int64)).cuda()
    pos0 = torch.tensor(z["points"][:, :2]).float().cuda()
    n = pos0.shape[0]
    torch.manual_seed(11)
    F, d, T = 5, 2, 7
    o0, o1, nt = 0, 2, 4
    model = EncodeProcessDecode(3, F, 3, d).cuda().eval()

    def predict(g):
        return g.x[:, o0:o1] + model(g)

    node_type = torch.tensor([0, 4, 5, 6])[torch.randint(0, 4, (n,))].float().cuda()
    extra = torch.randint(0, n, (2, 2000), device="cuda")
    extra = extra[:, extra[0] != extra[1]]
    ei_more = G.to_undirected(torch.cat([ei, extra], dim=1), n)          # frames 3 and 4: additional ("world") edges
    assert ei_more.shape[1] > ei.shape[1]
    frames = []
    for t in range(T):
        x_t = torch.randn(n, F, device="cuda")
        x_t[:, nt] = node_type
        pos_t = pos0 * (1.0 + 0.3 * t) + 0.05 * t * torch.randn(n, 2, device="cuda")
        top = ei_more if t in (3, 4) else ei
        frames.append((x_t, torch.randn(n, d, device="cuda"), G.edge_features(pos_t, top), top))
    graph = Graph(x=torch.zeros(n, F, device="cuda"), edge_attr=frames[0][2].clone(), edge_index=ei)
    want = _literal_loop(predict, graph, frames, o0, o1, nt, None)
    got = Rollout(predict, graph.clone(), (o0, o1), nt, capture=True)
    # the clone has its own edge_index object: frames must name THAT object to replay
    fr2 = [(x, y, ea, (got.graph.edge_index if top is ei else top)) for x, y, ea, top in frames]
    out = got.run(fr2)
    torch.cuda.synchronize()
    scale = float(want.abs().max())
    assert float((out - want).abs().max()) <= 1e-3 * scale
    # frame-0 edges for every frame (what round 1 silently did) is a different answer
    stale = Rollout(predict, graph.clone(), (o0, o1), nt, capture=True).run([(x, y) for x, y, _, _ in frames])
    assert float((stale - want).abs().max()) > 10 * float((out - want).abs().max()) + 1e-3 * scale
    # mismatching edge features raise instead of being ignored
    with pytest.raises(ValueError):
        got.step(frames[0][0], frames[0][1], edge_attr=frames[3][2])
    with pytest.raises(ValueError):
        got.step(frames[0][0], frames[0][1], edge_index=ei_more)
