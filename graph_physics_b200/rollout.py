"""Autoregressive rollout: the forward-only loop of the reference's validation / predict steps
(`LightningModule._make_prediction`, graphphysics/training/lightning_module.py:168-202, driven by `validation_step`
:204-232 and `predict_step` :300-329) with everything a trajectory shares kept on the device -- topology and its CSR,
packed weights, every buffer -- and ONE captured CUDA graph replayed per time step.

Reference semantics of one step (batch size 1, topology constant along the trajectory):

    batch = batch.clone()
    if last_prediction is not None:
        batch.x[:, out_start:out_end] = last_prediction
        if use_previous_data: batch.x[:, prev_start:prev_end] = last_previous_data_prediction
    mask = build_mask(param, batch)            # NOT (NORMAL or OUTFLOW): boundary nodes, from batch.x[:, node_type_index]
    current_output = batch.x[:, out_start:out_end]
    predicted = model(batch)                   # no_grad
    predicted[mask] = target[mask]             # boundary nodes reset to the ground truth
    last_prediction = predicted
    last_previous_data_prediction = predicted - current_output

Here the first step is made uniform with the others by initialising `last_prediction` / `last_previous_data_prediction`
from the first frame's own columns (overwriting a column with itself), so a single graph serves the whole trajectory.
A batch of a few hundred sequential steps on a ~2 k-node mesh is launch-latency bound when run eagerly (~150 kernel
launches of a few microseconds each per step); the replay removes that.

Moving meshes / world edges (DeformingPlate; the reference rebuilds FaceToEdge, `add_world_edges`, Cartesian / Distance
and the world-position features for EVERY frame, graphphysics/dataset/preprocessing.py:401-427): a frame may carry its
own `edge_attr` and `edge_index`.  Edge features of the captured topology are copied into the static buffer and the
replay continues; a different topology (another `edge_index` tensor) runs that step eagerly on the frame's own edges.
A frame whose `edge_attr` does not match the topology it is used with raises -- nothing falls back to frame-0 edges.

`predict_fn(graph) -> [N, d]` is any forward-only callable built on this package's modules (e.g. an
`EncodeProcessDecode` wrapped with the simulator's pre / post-processing); it must read `graph.x` (and
`graph.edge_attr`, `graph.edge_index`) from the tensors it is handed, because those are the static buffers of the capture.
"""
from __future__ import annotations

from typing import Callable, Iterable, Optional, Sequence, Tuple

import torch

from . import _lib as L

# reference graphphysics/utils/nodetype.py:4-12
NODE_NORMAL, NODE_OUTFLOW = 0, 5


def boundary_mask(x: torch.Tensor, node_type_index: int) -> torch.Tensor:
    """`build_mask` (lightning_module.py:17-25): True where the node is neither NORMAL nor OUTFLOW."""
    nt = x[:, 0, node_type_index] if x.dim() > 2 else x[:, node_type_index]
    return torch.logical_not(torch.logical_or(nt == NODE_NORMAL, nt == NODE_OUTFLOW))


class Rollout:
    """rollout = Rollout(predict_fn, graph, out_slice=(s, e), node_type_index=i, prev_slice=None)
    preds = rollout.run(frames)          # frames: iterable of (x_t [N,F], y_t [N,d][, edge_attr_t[, edge_index_t]])
                                         # -> [T,N,d] on the device

    `graph` is duck typed (`.x`, `.edge_attr`, `.edge_index`, optional `.pos`); its tensors become the static
    buffers.  `capture=False` runs the same step eagerly (the check the tests compare the replay against)."""

    def __init__(self, predict_fn: Callable, graph, out_slice: Tuple[int, int], node_type_index: int,
                 prev_slice: Optional[Tuple[int, int]] = None, capture: bool = True, warmup: int = 2):
        L.require_device(graph.x)
        self.predict_fn = predict_fn
        self.graph = graph
        self.o0, self.o1 = int(out_slice[0]), int(out_slice[1])
        self.prev = (int(prev_slice[0]), int(prev_slice[1])) if prev_slice is not None else None
        self.nt = int(node_type_index)
        n, d = graph.x.shape[0], self.o1 - self.o0
        dev = graph.x.device
        self.y = torch.zeros((n, d), dtype=graph.x.dtype, device=dev)
        self.last = torch.zeros((n, d), dtype=graph.x.dtype, device=dev)
        self.last_prev = torch.zeros((n, d), dtype=graph.x.dtype, device=dev) if self.prev is not None else None
        self.pred = torch.zeros((n, d), dtype=graph.x.dtype, device=dev)
        self.capture = bool(capture)
        self.warmup = int(warmup)
        self._graph = None
        # the topology the static buffers (and the capture) belong to
        self._ei0 = graph.edge_index
        self._ea0 = getattr(graph, "edge_attr", None)

    # one step on the static buffers (graph.x already holds frame t's features, self.y its target)
    def _step(self):
        x = self.graph.x
        x[:, self.o0:self.o1] = self.last
        if self.prev is not None:
            x[:, self.prev[0]:self.prev[1]] = self.last_prev
        mask = boundary_mask(x, self.nt)
        cur = x[:, self.o0:self.o1].clone()
        with torch.no_grad():
            out = self.predict_fn(self.graph)
        out = torch.where(mask.unsqueeze(-1), self.y, out)
        if self.prev is not None:
            self.last_prev.copy_(out - cur)
        self.last.copy_(out)
        self.pred.copy_(out)

    def _ensure_captured(self):
        if self._graph is not None or not self.capture:
            return
        keep = [t.clone() for t in (self.graph.x, self.last, self.pred)]
        keep_prev = self.last_prev.clone() if self.last_prev is not None else None
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup):
                self._step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self._step()
        self._graph = g
        # the warm-up / capture runs advanced the state: put it back
        self.graph.x.copy_(keep[0]); self.last.copy_(keep[1]); self.pred.copy_(keep[2])
        if keep_prev is not None:
            self.last_prev.copy_(keep_prev)

    def reset(self, x0: torch.Tensor):
        """Start of a trajectory (`_reset_validation_trajectory`, lightning_module.py:163-166): no previous prediction;
        the first step then keeps the first frame's own output / previous-data columns."""
        x0 = x0.to(self.graph.x.device)
        self.last.copy_(x0[:, self.o0:self.o1])
        if self.prev is not None:
            self.last_prev.copy_(x0[:, self.prev[0]:self.prev[1]])

    def step(self, x_t: torch.Tensor, y_t: torch.Tensor, edge_attr: Optional[torch.Tensor] = None,
             edge_index: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One `_make_prediction` call: frame features and target in, prediction out (a view of a static buffer:
        copy it if it must survive the next step).
        edge_attr: this frame's edge features (moving mesh); edge_index: this frame's topology if it is not the one the
        rollout was built on -- pass the SAME tensor object (or None) for frames that keep it, so that they replay."""
        self.graph.x.copy_(x_t, non_blocking=True)
        self.y.copy_(y_t, non_blocking=True)
        same_topology = edge_index is None or edge_index is self._ei0
        if same_topology:
            if edge_attr is not None:
                if self._ea0 is None or tuple(edge_attr.shape) != tuple(self._ea0.shape):
                    raise ValueError(f"rollout: edge_attr {tuple(edge_attr.shape)} does not match the topology the "
                                     f"rollout was built on ({None if self._ea0 is None else tuple(self._ea0.shape)}); "
                                     "pass this frame's edge_index as well")
                self._ea0.copy_(edge_attr, non_blocking=True)
            self.graph.edge_index, self.graph.edge_attr = self._ei0, self._ea0
            if self.capture:
                self._ensure_captured()
                self._graph.replay()
            else:
                self._step()
            return self.pred
        # another topology: this step runs eagerly on the frame's own edges (the capture belongs to self._ei0)
        if edge_index.dim() != 2 or edge_index.shape[0] != 2:
            raise ValueError("rollout: edge_index must be [2, E]")
        if self._ea0 is not None:
            if edge_attr is None or edge_attr.shape[0] != edge_index.shape[1] or edge_attr.shape[1] != self._ea0.shape[1]:
                raise ValueError("rollout: a frame with its own edge_index needs its own edge_attr [E, "
                                 f"{self._ea0.shape[1]}]")
        self.graph.edge_index, self.graph.edge_attr = edge_index, edge_attr
        try:
            self._step()
        finally:
            self.graph.edge_index, self.graph.edge_attr = self._ei0, self._ea0
        return self.pred

    def run(self, frames: Iterable[Sequence[torch.Tensor]]) -> torch.Tensor:
        out = []
        first = True
        for fr in frames:
            x_t, y_t = fr[0], fr[1]
            ea_t = fr[2] if len(fr) > 2 else None
            ei_t = fr[3] if len(fr) > 3 else None
            if first:
                self.graph.x.copy_(x_t, non_blocking=True)      # (the capture warm-up must see real features)
                self.reset(x_t)
                first = False
            out.append(self.step(x_t, y_t, ea_t, ei_t).clone())
        return torch.stack(out) if out else torch.empty((0,) + tuple(self.pred.shape), device=self.pred.device)
