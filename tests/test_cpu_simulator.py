"""CPU: the steps either side of the processor -- `Simulator`, `Normalizer`, masked losses --
against tests/golden/ref_simulator.npz, which tests/golden/make_golden_simulator.py wrote by running the reference's own
modules (graphphysics/models/simulator.py, models/layers.py:265-393, utils/loss.py:10-112), and the cases of the
reference's own tests (tests/graphphysics/models/test_simulator.py, test_layers.py:107-115).

These are fixed sequences of torch operations (no kernels of this package), so they run here; the fp32 tolerance covers
the different summation order of the masked mean."""
import os

import numpy as np
import pytest
import torch

from graph_physics_b200.models import Graph, Normalizer, Simulator
from graph_physics_b200.utils import L1SmoothLoss, L2Loss, NodeType

HERE = os.path.dirname(os.path.abspath(__file__))


def _gen():
    """The seeded input generator of the fixture script (the script's reference-reading part is not executed)."""
    src = open(os.path.join(HERE, "golden", "make_golden_simulator.py")).read()
    start, stop = src.index("N_NODES, N_EDGES"), src.index("def main():")
    ns = {"torch": torch}
    exec(compile(src[start:stop], "make_golden_simulator_inputs", "exec"), ns)
    return ns


G = _gen()
F_IN, E_IN, D_OUT, NT_IDX = G["F_IN"], G["E_IN"], G["D_OUT"], G["NODE_TYPE_INDEX"]


class _Net(torch.nn.Module):
    K = 0

    def __init__(self):
        super().__init__()
        self.w, self.b = G["net_weight"]()

    def forward(self, graph):
        return graph.x @ self.w + self.b + 0.01 * graph.edge_attr.sum()


def _simulator(model=None):
    return Simulator(node_input_size=F_IN + NodeType.SIZE, edge_input_size=E_IN, output_size=D_OUT,
                     feature_index_start=0, feature_index_end=F_IN, output_index_start=0, output_index_end=D_OUT,
                     node_type_index=NT_IDX, model=model or _Net(), device=torch.device("cpu"))


def _close(a, b, tol=2e-6):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape
    assert np.abs(a - b).max() <= tol * max(1.0, np.abs(b).max()), np.abs(a - b).max()


def test_simulator_matches_reference_fixture(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_simulator.npz"))
    sim = _simulator()
    net = sim.model
    masks = [NodeType.NORMAL, NodeType.OUTFLOW]
    sim.train()
    for c in range(G["N_CALLS"]):
        x, y, ea, ei, pos = G["inputs"](c)
        graph, tdn = sim._build_input_graph(Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos), is_training=True)
        _close(graph.x, z[f"train{c}_node_features"])
        _close(graph.edge_attr, z[f"train{c}_edge_attr"])
        _close(tdn, z[f"train{c}_target"])
        assert graph.edge_index is ei and graph.pos is pos
        out = net(graph)
        _close(L2Loss()(tdn, out, x[:, NT_IDX], masks=masks), z[f"train{c}_l2"])
        _close(L1SmoothLoss(beta=0.5)(tdn, out, x[:, NT_IDX], masks=masks), z[f"train{c}_l1s"])
    sim.eval()
    x, y, ea, ei, pos = G["inputs"](G["N_CALLS"])
    net_out, tdn, outputs = sim(Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos))
    _close(net_out, z["eval_network_output"])
    _close(tdn, z["eval_target"])
    _close(outputs, z["eval_outputs"])
    for nme in ("_output_normalizer", "_node_normalizer", "_edge_normalizer"):
        for k in ("_acc_count", "_num_accumulations", "_acc_sum", "_acc_sum_squared"):
            _close(getattr(getattr(sim, nme), k), z[f"{nme}.{k}"], tol=1e-6)


def test_simulator_shapes_as_in_the_reference_tests():
    """tests/graphphysics/models/test_simulator.py:62-113 (10 nodes, 15 edges, mock network)."""
    class Mock(torch.nn.Module):
        K = 0

        def forward(self, graph):
            return torch.randn(graph.x.size(0), 2)

    sim = _simulator(Mock())
    x = torch.randn(10, 6)
    x[:, 5] = torch.abs(x[:, 5])
    data = Graph(x=x, y=torch.randn(10, 2), edge_attr=torch.randn(15, 4), edge_index=torch.randint(0, 10, (2, 15)),
                 pos=torch.randn(10, 3))
    sim.train()
    out, tdn, outputs = sim(data)
    assert outputs is None and out.shape == (10, 2) and tdn.shape == (10, 2)
    sim.eval()
    out, tdn, outputs = sim(data)
    assert outputs.shape == (10, 2) and out.shape == (10, 2) and tdn.shape == (10, 2)
    assert sim._get_pre_target(data).shape == (10, 2)
    one_hot = sim._get_one_hot_type(data)
    assert one_hot.shape == (10, NodeType.SIZE)
    assert torch.equal(one_hot, torch.nn.functional.one_hot(x[:, 5].long(), NodeType.SIZE).float())
    assert sim._build_node_features(data, one_hot).shape == (10, 5 + NodeType.SIZE)
    assert sim._build_outputs(data, torch.randn(10, 2)).shape == (10, 2)
    # no edge features: no edge normaliser, edge_attr handed through (simulator.py:165, 194-198, 290-293)
    sim0 = Simulator(5 + NodeType.SIZE, 0, 2, 0, 5, 0, 2, 5, Mock(), torch.device("cpu"))
    assert sim0._edge_normalizer is None and sim0.edge_input_size is None
    graph, _ = sim0._build_input_graph(data, is_training=False)
    assert graph.edge_attr is data.edge_attr


def test_normalizer_inverse_and_accumulation_gate():
    """test_layers.py:107-115, plus the `max_accumulations` stop (layers.py:329-332) decided on the device."""
    nz = Normalizer(7, device="cpu")
    x = torch.randn(10, 7)
    y = nz(x)
    assert y.shape == x.shape
    assert torch.allclose(nz.inverse(y), x, atol=1e-6)
    assert abs(float(y.mean())) < 1e-6 and abs(float(y.std(unbiased=False)) - 1.0) < 1e-2
    nz = Normalizer(3, max_accumulations=2, device="cpu")
    a, b, c = torch.randn(5, 3), torch.randn(6, 3) + 1.0, torch.randn(50, 3) * 9.0
    nz(a)
    nz(b)
    before = {k: v.clone() for k, v in nz.state_dict().items()}
    nz(c)                                   # third batch: not accumulated
    for k, v in nz.state_dict().items():
        assert torch.equal(v, before[k]), k
    assert float(nz._acc_count) == 11.0 and float(nz._num_accumulations) == 2.0
    ab = torch.cat([a, b])
    assert torch.allclose(nz._mean(), ab.mean(0, keepdim=True), atol=1e-6)
    assert torch.allclose(nz._std_with_epsilon(), ab.std(0, unbiased=False, keepdim=True), atol=1e-5)
    nz(c, accumulate=False)
    assert float(nz._acc_count) == 11.0
    # an empty normaliser: mean 0, std = epsilon (layers.py:363-375)
    nz = Normalizer(2, device="cpu")
    assert torch.equal(nz._mean(), torch.zeros(1, 2)) and float(nz._std_with_epsilon().max()) == pytest.approx(1e-8)
    assert set(nz.get_variable()) == {"_max_accumulations", "_std_epsilon", "_acc_count", "_num_accumulations",
                                      "_acc_sum", "_acc_sum_squared", "name"}


def test_masked_losses_equal_the_boolean_index_formulation():
    """utils/loss.py:10-112 literally (errors[mask] then mean) against the fixed-shape masked mean used here, with and
    without `selected_indexes`; non-finite errors on excluded nodes must not leak in."""
    g = torch.Generator().manual_seed(5)
    n, d = 203, 3
    target, out = torch.randn(n, d, generator=g), torch.randn(n, d, generator=g)
    node_type = torch.randint(0, 7, (n,), generator=g).float()
    sel = torch.randperm(n, generator=g)[:40]
    masks = [NodeType.NORMAL, NodeType.OUTFLOW]
    for selected in (None, sel):
        m = (node_type == 0) | (node_type == 5)
        if selected is not None:
            m = m & ~torch.isin(torch.arange(n), selected)
        want_l2 = ((out - target) ** 2)[m].mean()
        want_l1 = torch.nn.functional.smooth_l1_loss(out, target, reduction="none", beta=0.7)[m].mean()
        assert torch.allclose(L2Loss()(target, out, node_type, masks, selected), want_l2, rtol=1e-6)
        assert torch.allclose(L1SmoothLoss(beta=0.7)(target, out, node_type, masks, selected), want_l1, rtol=1e-6)
    bad = out.clone()
    bad[node_type == 3] = float("inf")
    assert torch.isfinite(L2Loss()(target, bad, node_type, masks))
    assert torch.isnan(L2Loss()(target, out, node_type, [NodeType.SIZE]))          # empty selection, as torch.mean([])
    assert L2Loss().__name__ == "MSE" and L1SmoothLoss().__name__ == "L1Smooth"
    # gradient: d loss / d out = 2 (out - target) / (count * d) on the selected rows, 0 elsewhere
    o = out.clone().requires_grad_(True)
    L2Loss()(target, o, node_type, masks).backward()
    m = (node_type == 0) | (node_type == 5)
    want = torch.where(m.unsqueeze(-1), 2 * (out - target) / (m.sum() * d), torch.zeros(()))
    assert torch.allclose(o.grad, want, atol=1e-7)


def test_checkpoint_round_trip_keeps_reference_layout(tmp_path):
    """simulator.py:359-405: {"model": state_dict, "_output_normalizer": {...}, "_node_normalizer", "_edge_normalizer"}."""
    sim = _simulator()
    sim.train()
    x, y, ea, ei, pos = G["inputs"](0)
    sim(Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos))
    path = os.path.join(tmp_path, "ckpt", "simulator.pth")
    sim.save_checkpoint(path)
    blob = torch.load(path, weights_only=False)
    assert set(blob) == {"model", "_output_normalizer", "_node_normalizer", "_edge_normalizer"}
    assert "_node_normalizer._acc_sum" in blob["model"] and "_output_normalizer._acc_count" in blob["model"]
    other = _simulator()
    ptr = other._node_normalizer._acc_sum.data_ptr()
    other.load_checkpoint(path)
    assert other._node_normalizer._acc_sum.data_ptr() == ptr          # buffers updated in place
    for k, v in sim.state_dict().items():
        assert torch.equal(v, other.state_dict()[k]), k
    sim.eval(), other.eval()
    a = sim(Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos))
    b = other(Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos))
    assert all(torch.equal(u, v) for u, v in zip(a, b))
    sim.freeze_all()
    assert not any(p.requires_grad for p in sim.model.parameters())


def test_predict_equals_the_evaluation_outputs_without_a_target():
    sim = _simulator()
    sim.train()
    for c in range(2):
        x, y, ea, ei, pos = G["inputs"](c)
        sim(Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos))
    sim.eval()
    x, y, ea, ei, pos = G["inputs"](2)
    before = {k: v.clone() for k, v in sim.state_dict().items()}
    _, _, outputs = sim(Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos))
    got = sim.predict(Graph(x=x, edge_attr=ea, edge_index=ei, pos=pos))          # no y
    assert torch.equal(got, outputs)
    sim.train()
    assert torch.equal(sim.predict(Graph(x=x, edge_attr=ea, edge_index=ei)), outputs)   # never accumulates
    for k, v in sim.state_dict().items():
        assert torch.equal(v, before[k]), k


def test_training_step_function_matches_the_reference_loss_fixture(golden_dir):
    """training/lightning_module.py:112-123 as a function: the fixture's L2 losses (reference Simulator + L2Loss on the
    same three accumulating calls), and the gradient reaches the network."""
    from graph_physics_b200.training import rollout_rmse, training_step
    z = np.load(os.path.join(golden_dir, "ref_simulator.npz"))

    class Net(torch.nn.Module):
        K = 0

        def __init__(self):
            super().__init__()
            w, b = G["net_weight"]()
            self.w, self.b = torch.nn.Parameter(w), torch.nn.Parameter(b)

        def forward(self, graph):
            return graph.x @ self.w + self.b + 0.01 * graph.edge_attr.sum()

    sim = _simulator(Net()).train()
    for c in range(G["N_CALLS"]):
        x, y, ea, ei, pos = G["inputs"](c)
        loss = training_step(sim, Graph(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos), L2Loss(),
                             [NodeType.NORMAL, NodeType.OUTFLOW])
        _close(loss.detach(), z[f"train{c}_l2"])
    loss.backward()
    assert sim.model.w.grad is not None and float(sim.model.w.grad.abs().sum()) > 0
    pred = torch.zeros(3, 5, 2)
    tgt = [torch.full((5, 2), float(i)) for i in range(3)]
    assert torch.allclose(rollout_rmse(pred, tgt), torch.tensor([0.0, 1.0, 2.0]))
