#!/usr/bin/env python
"""Fixture for the steps either side of the processor: runs the REFERENCE's own `Simulator`
(graphphysics/models/simulator.py:128-342), `Normalizer` (models/layers.py:265-393) and `L2Loss` / `L1SmoothLoss`
(utils/loss.py:10-112) from /root/reference on seeded inputs and writes tests/golden/ref_simulator.npz.

    python tests/golden/make_golden_simulator.py

torch_geometric is not installed: the stand-ins of make_golden.py are registered first (only `Data` is touched on
this path).  The network inside the simulator is a fixed linear map of the normalised node features (K = 0), so the
fixture pins exactly the pre / post-processing: target delta, one-hot node types, the three running normalisers over
several accumulating training calls, the evaluation-mode reconstruction `pre_target + inverse(output)` and the
masked losses.  Build container only (reads /root/reference); the tests use the committed output.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, _install_stubs  # noqa: E402

N_NODES, N_EDGES, N_CALLS = 97, 311, 3
F_IN, E_IN, D_OUT = 5, 4, 2          # raw node features [v(2) | extra(3) | node_type], edge features, output width
NODE_TYPE_INDEX = 5


def inputs(call: int):
    g = torch.Generator().manual_seed(1000 + call)
    x = torch.randn(N_NODES, F_IN + 1, generator=g) * (1.0 + call) + 0.3 * call
    x[:, NODE_TYPE_INDEX] = torch.randint(0, 7, (N_NODES,), generator=g).float()
    y = x[:, :D_OUT] + 0.1 * torch.randn(N_NODES, D_OUT, generator=g)
    edge_attr = torch.randn(N_EDGES, E_IN, generator=g) * 2.0 - 0.5
    edge_index = torch.randint(0, N_NODES, (2, N_EDGES), generator=g)
    pos = torch.randn(N_NODES, 2, generator=g)
    return x, y, edge_attr, edge_index, pos


def net_weight():
    g = torch.Generator().manual_seed(77)
    return torch.randn(F_IN + 9, D_OUT, generator=g) * 0.2, torch.randn(D_OUT, generator=g) * 0.1


def main():
    Data = _install_stubs()
    sys.path.insert(0, REF)
    from graphphysics.models.simulator import Simulator
    from graphphysics.utils.loss import L1SmoothLoss, L2Loss
    from graphphysics.utils.nodetype import NodeType

    w, b = net_weight()

    class Net(torch.nn.Module):
        K = 0

        def forward(self, graph):
            return graph.x @ w + b + 0.01 * graph.edge_attr.sum()

    sim = Simulator(node_input_size=F_IN + NodeType.SIZE, edge_input_size=E_IN, output_size=D_OUT,
                    feature_index_start=0, feature_index_end=F_IN, output_index_start=0, output_index_end=D_OUT,
                    node_type_index=NODE_TYPE_INDEX, model=Net(), device=torch.device("cpu"))
    masks = [NodeType.NORMAL, NodeType.OUTFLOW]
    out = {}
    sim.train()
    for c in range(N_CALLS):
        x, y, ea, ei, pos = inputs(c)
        d = Data(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos)
        graph, tdn = sim._build_input_graph(d, is_training=True)       # accumulates once
        sim2_out = graph.x @ w + b + 0.01 * graph.edge_attr.sum()
        out[f"train{c}_node_features"] = graph.x.numpy()
        out[f"train{c}_edge_attr"] = graph.edge_attr.numpy()
        out[f"train{c}_target"] = tdn.numpy()
        out[f"train{c}_l2"] = L2Loss()(tdn, sim2_out, x[:, NODE_TYPE_INDEX], masks=masks).numpy()
        out[f"train{c}_l1s"] = L1SmoothLoss(beta=0.5)(tdn, sim2_out, x[:, NODE_TYPE_INDEX], masks=masks).numpy()
    sim.eval()
    x, y, ea, ei, pos = inputs(N_CALLS)
    net_out, tdn, outputs = sim(Data(x=x, y=y, edge_attr=ea, edge_index=ei, pos=pos))
    out["eval_network_output"] = net_out.numpy()
    out["eval_target"] = tdn.numpy()
    out["eval_outputs"] = outputs.numpy()
    for nme in ("_output_normalizer", "_node_normalizer", "_edge_normalizer"):
        nz = getattr(sim, nme)
        for k in ("_acc_count", "_num_accumulations", "_acc_sum", "_acc_sum_squared"):
            out[f"{nme}.{k}"] = getattr(nz, k).numpy()
    np.savez_compressed(os.path.join(HERE, "ref_simulator.npz"), **out)
    print("wrote ref_simulator.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
