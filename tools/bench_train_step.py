#!/usr/bin/env python
"""`training_step` of the reference (training/lightning_module.py:112-123) end to end on the device: raw node features
with a node-type column -> `Simulator` (target delta, one-hot, three accumulating normalisers) -> EncodeProcessDecode
(15 blocks) -> masked `L2Loss` -> backward, 32 x cylinder mesh.  Times the eager step and the same step replayed from ONE
captured CUDA graph (possible because the Simulator / Normalizer / loss of this package read nothing back on the host).

    python tools/bench_train_step.py            # one JSON line (needs a B200)

Not run yet: written after round 1's GPU budget was spent."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200 import graph as G, processor as proc  # noqa: E402
from graph_physics_b200.models import EncodeProcessDecode, Graph, Simulator  # noqa: E402
from graph_physics_b200.training import training_step  # noqa: E402
from graph_physics_b200.utils import L2Loss, NodeType  # noqa: E402

z = np.load(os.path.join(ROOT, "tests", "golden", "cylinder_mesh.npz"))
n1 = z["points"].shape[0]
copies = 32
dev = torch.device("cuda")
ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).to(dev), n1, copies)
pos = torch.as_tensor(z["points"][:, :2]).to(dev).repeat(copies, 1)
N, E = n1 * copies, ei.shape[1]
torch.manual_seed(0)
# CylinderFlow layout: x = [velocity(2) | node_type], node features = velocity + one-hot(9) = 11, edge features 3
x = torch.randn(N, 3, device=dev)
x[:, 2] = torch.randint(0, 7, (N,), device=dev).float()
y = x[:, :2] + 0.1 * torch.randn(N, 2, device=dev)
ea = G.edge_features(pos, ei)
model = EncodeProcessDecode(15, 2 + NodeType.SIZE, 3, 2).to(dev)
sim = Simulator(node_input_size=2 + NodeType.SIZE, edge_input_size=3, output_size=2, feature_index_start=0,
                feature_index_end=2, output_index_start=0, output_index_end=2, node_type_index=2, model=model,
                device=dev).train()
loss_fn = L2Loss()
masks = [NodeType.NORMAL, NodeType.OUTFLOW]
params = [p for p in sim.parameters() if p.requires_grad]
batch = Graph(x=x, y=y, pos=pos, edge_attr=ea, edge_index=ei)
loss_out = torch.zeros((), device=dev)


def step():
    for p in params:
        p.grad = None
    loss = training_step(sim, batch, loss_fn, masks)
    loss.backward()
    loss_out.copy_(loss.detach())


def timed(fn, steps=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / steps


res = {"eager": timed(step)}
cap = proc.CapturedStep(step)
res["captured"] = timed(cap)
print(json.dumps({"workload": f"training_step: Simulator + EncodeProcessDecode(15, 11, 3, 2) + masked L2 loss, fwd+bwd, "
                              f"32 x cylinder mesh, N={N}, E={E}", "ms_per_step": res, "loss": float(loss_out),
                  "edges_per_s_captured": E / (res["captured"] * 1e-3)}))
