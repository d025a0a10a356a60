#!/usr/bin/env python
"""Whole-model step (SURVEY.md 8(d) "also report": full training_step for the CylinderFlow config): encoders +
15 processor blocks + decoder, forward + backward of an MSE loss, 32 x cylinder mesh (N=61,536, E=354,240,
node inputs 2+9, edge inputs 3, outputs 2).  Times the step with the encoders on the fused node kernels
(graph_physics_b200.encoder) and with eager torch encoders, both replayed from a CUDA graph."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200 import graph as G, processor as proc  # noqa: E402
from graph_physics_b200.models import EncodeProcessDecode, Graph  # noqa: E402

z = np.load(os.path.join(ROOT, "tests", "golden", "cylinder_mesh.npz"))
n1 = z["points"].shape[0]
copies = 32
ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, copies)
N, E = n1 * copies, ei.shape[1]
torch.manual_seed(0)
model = EncodeProcessDecode(15, 11, 3, 2).cuda()
x = torch.randn(N, 11, device="cuda")
ea = torch.randn(E, 3, device="cuda")
y = torch.randn(N, 2, device="cuda")
params = list(model.parameters())


def step():
    for p in params:
        p.grad = None
    out = model(Graph(x=x, edge_index=ei, edge_attr=ea))
    loss = (out - y).square().mean()
    loss.backward()
    return loss


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


res = {}
for fused in (True, False):
    model.fused_encoders = fused
    cap = proc.CapturedStep(step)
    res["fused_encoders" if fused else "eager_encoders"] = timed(cap)
    del cap
only = EncodeProcessDecode(15, 128, 128, 128, only_processor=True).cuda()
only.load_state_dict({k: v for k, v in model.state_dict().items() if k.startswith("processor_list")})
xl, el, gl = torch.randn(N, 128, device="cuda"), torch.randn(E, 128, device="cuda"), torch.randn(N, 128, device="cuda")
oparams = list(only.parameters())


def pstep():
    for p in oparams:
        p.grad = None
    only(Graph(x=xl, edge_index=ei, edge_attr=el)).backward(gl)


cap = proc.CapturedStep(pstep)
res["processor_only"] = timed(cap)
print(json.dumps({"workload": f"EncodeProcessDecode(15, 11, 3, 2), 32 x cylinder mesh, N={N}, E={E}, fwd+bwd of an MSE "
                              "loss, CUDA-graph replay", "ms_per_step": res,
                  "edges_per_s_full_model": E / (res["fused_encoders"] * 1e-3)}))
