"""ORACLE / TEST INFRASTRUCTURE ONLY -- CPU restatement (plain torch ops, fp32 or
fp64) of the reference's MeshGraphNet processor (SURVEY.md section 8(a) rows
A1-A8; backward B1-B3 comes from autograd over these ops).

Only tests/, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl
reference` legs may import this file.  The product path (graph_physics_b200/)
never does and has no CPU fallback.

Parity status
-------------
The reference's own tests for this path assert shapes / no-NaN / grad-exists only
(tests/graphphysics/models/test_layers.py:180-241, test_processors.py:22-126), so
floating point is "parity unpinned" BY THE REFERENCE'S TEST SUITE.  It is pinned
here instead by executing the reference's own `graphphysics/models/layers.py` and
`processors.py` source in the build container (tests/golden/make_golden.py, with a
minimal stand-in for the un-vendored torch_geometric 2.6.1 `MessagePassing`
scatter-sum) and committing its outputs as tests/golden/ref_*.npz; the test
tests/test_oracle_golden.py requires this file to reproduce them to 1e-6.

Functional style: parameters are passed as a flat dict with the reference's
state-dict keys (`processor_list.{k}.edge_block.{0,2,4,6}.{weight,bias}`,
`...edge_block.7.scale`, same for `node_block`, `nodes_encoder`, `edges_encoder`,
`decode_module`).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch


def rms_norm(x: torch.Tensor, scale: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    """RMSNorm.forward with p=-1, bias=False (graphphysics/models/layers.py:49-74):
    y = scale * x / (||x||_2 * d^-1/2 + eps); eps is added to the RMS itself."""
    d = x.shape[-1]
    rms = x.norm(2, dim=-1, keepdim=True) * d ** (-0.5)
    return scale * (x / (rms + eps))


def r16(t: torch.Tensor) -> torch.Tensor:
    """fp16 rounding with a straight-through gradient.  Used only by the `emulate_fp16`
    variants below, which restate WHERE the sm_100a kernels round their tensor-core operands
    (weights, e, x, aggr, P and the hidden activations are fp16 operands; accumulation,
    bias, RMSNorm, the segmented sum and both residual streams are fp32)."""
    return t + (t.detach().half().to(t.dtype) - t.detach())


def mlp(x: torch.Tensor, p: Dict[str, torch.Tensor], prefix: str,
        nb_of_layers: int = 4, layer_norm: bool = True, emulate_fp16: bool = False,
        first_layer_extra: Optional[torch.Tensor] = None, masks=None) -> torch.Tensor:
    """build_mlp (layers.py:77-113): Linear,ReLU,(Linear,ReLU)*(L-2),Linear,[RMSNorm];
    Sequential indices 0,2,4,... are the Linear layers, 2L-1 the norm.

    emulate_fp16: operands of every Linear are rounded to fp16 (straight-through).
    first_layer_extra: the kernels' split of edge layer 1 -- the first Linear sees only the
    edge-latent columns of W1 and the per-node partial products are added (emulation only).
    masks: optional list of 0/1 tensors; ReLU i is replaced by multiplication with masks[i]
    (lets a test differentiate THROUGH THE SAME ACTIVE SET as the device kernels, whose
    fp32-accumulated pre-activations can differ in sign from fp64 ones when they are ~0)."""
    h = x
    for i in range(nb_of_layers):
        w = p[f"{prefix}.{2 * i}.weight"]
        b = p[f"{prefix}.{2 * i}.bias"]
        if emulate_fp16:
            w = r16(w)
            h = r16(h)
        if i == 0 and first_layer_extra is not None:
            h = torch.nn.functional.linear(h, w[:, :h.shape[-1]]) + first_layer_extra
        else:
            h = torch.nn.functional.linear(h, w, b)
        if i < nb_of_layers - 1:
            h = torch.relu(h) if masks is None else h * masks[i]
    if layer_norm:
        h = rms_norm(h, p[f"{prefix}.{2 * nb_of_layers - 1}.scale"])
    return h


def graph_net_block(x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor,
                    p: Dict[str, torch.Tensor], prefix: str = "",
                    nb_of_layers: int = 4, layer_norm: bool = True, emulate_fp16: bool = False,
                    edge_masks=None, node_masks=None):
    """GraphNetBlock.forward (layers.py:667-701).

    row = edge_index[0] (sender j), col = edge_index[1] (receiver i);
    e_ = edge_block(cat[e, x[col], x[row]])           (layers.py:687-691, 717-719)
    a  = sum over edges of e_ at index col, size N     (layers.py:694-696; PyG
         MessagePassing(aggr="add", flow="source_to_target") -> scatter-sum)
    x_ = node_block(cat[x, a])                         (layers.py:744-746)
    returns (x + x_, e + e_)                           (layers.py:698-699)

    emulate_fp16=True restates the kernels' rounding points (see r16): the same function
    of the same parameters, with fp16 tensor-core operands and the algebraic split of edge
    layer 1 into per-node partial products P = [x W_i^T + b1 | x W_j^T] stored in fp16.
    """
    pre = prefix + "." if prefix and not prefix.endswith(".") else prefix
    row, col = edge_index[0], edge_index[1]
    if emulate_fp16:
        hd = x.shape[-1]
        w1 = r16(p[pre + "edge_block.0.weight"])
        x16 = r16(x)
        p_i = r16(torch.nn.functional.linear(x16, w1[:, hd:2 * hd], p[pre + "edge_block.0.bias"]))
        p_j = r16(torch.nn.functional.linear(x16, w1[:, 2 * hd:3 * hd]))
        extra = p_i.index_select(0, col) + p_j.index_select(0, row)
        e_ = mlp(edge_attr, p, pre + "edge_block", nb_of_layers, layer_norm, True, first_layer_extra=extra,
                 masks=edge_masks)
        aggr = torch.zeros_like(x).index_add_(0, col, e_)
        x_ = mlp(torch.cat([x, aggr], dim=-1), p, pre + "node_block", nb_of_layers, layer_norm, True,
                 masks=node_masks)
        return x + x_, edge_attr + e_
    x_i = x.index_select(0, col)
    x_j = x.index_select(0, row)
    e_ = mlp(torch.cat([edge_attr, x_i, x_j], dim=-1), p, pre + "edge_block", nb_of_layers, layer_norm)
    aggr = torch.zeros_like(x).index_add_(0, col, e_)
    x_ = mlp(torch.cat([x, aggr], dim=-1), p, pre + "node_block", nb_of_layers, layer_norm)
    return x + x_, edge_attr + e_


def processor(x: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor,
              p: Dict[str, torch.Tensor], num_steps: int, prefix: str = "processor_list",
              return_all: bool = False, emulate_fp16: bool = False):
    """The hot loop `for block in self.processor_list` (processors.py:130-131)."""
    xs, es = [], []
    for k in range(num_steps):
        x, edge_attr = graph_net_block(x, edge_index, edge_attr, p, f"{prefix}.{k}", emulate_fp16=emulate_fp16)
        if return_all:
            xs.append(x)
            es.append(edge_attr)
    if return_all:
        return x, edge_attr, xs, es
    return x, edge_attr


def encode_process_decode(graph_x: torch.Tensor, edge_index: torch.Tensor,
                          graph_edge_attr: torch.Tensor, p: Dict[str, torch.Tensor],
                          num_steps: int, only_processor: bool = False) -> torch.Tensor:
    """EncodeProcessDecode.forward (processors.py:111-137), plain decoder
    (num_mixture_components == 0): decoder MLP has no norm (processors.py:84-90)."""
    if only_processor:
        x, e = graph_x, graph_edge_attr
    else:
        x = mlp(graph_x, p, "nodes_encoder")
        e = mlp(graph_edge_attr, p, "edges_encoder")
    x, e = processor(x, edge_index, e, p, num_steps)
    if only_processor:
        return x
    return mlp(x, p, "decode_module", layer_norm=False)


# ----------------------------------------------------------------- parameters
def _linear_init(out_f: int, in_f: int, gen: torch.Generator, dtype):
    """torch.nn.Linear default init: kaiming_uniform(a=sqrt(5)) == U(+-1/sqrt(in))
    for the weight and U(+-1/sqrt(in)) for the bias."""
    bound = 1.0 / math.sqrt(in_f)
    w = (torch.rand(out_f, in_f, generator=gen, dtype=torch.float64) * 2 - 1) * bound
    b = (torch.rand(out_f, generator=gen, dtype=torch.float64) * 2 - 1) * bound
    return w.to(dtype), b.to(dtype)


def init_mlp_params(p, prefix, in_size, hidden, out_size, gen, dtype=torch.float32,
                    nb_of_layers=4, layer_norm=True):
    sizes = [in_size] + [hidden] * (nb_of_layers - 1) + [out_size]
    for i in range(nb_of_layers):
        w, b = _linear_init(sizes[i + 1], sizes[i], gen, dtype)
        p[f"{prefix}.{2 * i}.weight"] = w
        p[f"{prefix}.{2 * i}.bias"] = b
    if layer_norm:
        p[f"{prefix}.{2 * nb_of_layers - 1}.scale"] = torch.ones(out_size, dtype=dtype)


def init_params(num_steps: int, hidden: int = 128, node_in: Optional[int] = None,
                edge_in: Optional[int] = None, out_size: Optional[int] = None,
                seed: int = 0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Random parameters with the reference's state-dict keys and default-init
    distributions (layers.py:101-113 uses stock nn.Linear / ones for scale)."""
    gen = torch.Generator().manual_seed(seed)
    p: Dict[str, torch.Tensor] = {}
    if node_in is not None:
        init_mlp_params(p, "nodes_encoder", node_in, hidden, hidden, gen, dtype)
        init_mlp_params(p, "edges_encoder", edge_in, hidden, hidden, gen, dtype)
        init_mlp_params(p, "decode_module", hidden, hidden, out_size, gen, dtype, layer_norm=False)
    for k in range(num_steps):
        init_mlp_params(p, f"processor_list.{k}.edge_block", 3 * hidden, hidden, hidden, gen, dtype)
        init_mlp_params(p, f"processor_list.{k}.node_block", 2 * hidden, hidden, hidden, gen, dtype)
    return p


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    """||a-b||_2 / ||b||_2 in float64 (the tolerance metric of BASELINE.json)."""
    a64, b64 = a.detach().double().flatten(), b.detach().double().flatten()
    return float((a64 - b64).norm() / b64.norm().clamp_min(1e-300))
