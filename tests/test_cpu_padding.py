"""CPU: the zero-padding / identity-layer embedding that lets hidden_size < 128 and nb_of_layers < 4 blocks run on the
128-wide four-layer kernels (graph_physics_b200.models.layers.block_param_list) is EXACT algebra: a plain-torch
evaluation of the embedded block (with RMSNorm dividing by sqrt(d) of the true width, the kernels' `norm_dim`) equals
the oracle's evaluation of the original block, forward and -- through autograd of the padding ops -- backward.
Reference: graphphysics/models/layers.py:77-113 (build_mlp with nb_of_layers), :630-665 (GraphNetBlock ctor)."""
import pytest
import torch

from oracle import mgn_oracle as mo


def _mlp4(h, prm, d):
    w1, b1, w2, b2, w3, b3, w4, b4, sc = prm
    h = torch.relu(torch.nn.functional.linear(h, w1, b1))
    h = torch.relu(torch.nn.functional.linear(h, w2, b2))
    h = torch.relu(torch.nn.functional.linear(h, w3, b3))
    z = torch.nn.functional.linear(h, w4, b4)
    return sc * (z / (z.norm(2, dim=-1, keepdim=True) * d ** -0.5 + 1e-8))


def _embedded_block(x, ei, e, prm18, d):
    """What the kernels compute, in plain torch, on 128-wide padded latents."""
    row, col = ei
    e_ = _mlp4(torch.cat([e, x[col], x[row]], dim=-1), prm18[:9], d)
    aggr = torch.zeros_like(x).index_add_(0, col, e_)
    x_ = _mlp4(torch.cat([x, aggr], dim=-1), prm18[9:], d)
    return x + x_, e + e_


@pytest.mark.parametrize("hidden,layers", [(32, 2), (32, 3), (32, 4), (64, 4), (128, 2), (128, 3), (40, 3)])
def test_embedding_is_exact(hidden, layers):
    from graph_physics_b200.models.layers import GraphNetBlock, block_param_list
    torch.manual_seed(hidden * 10 + layers)
    blk = GraphNetBlock(hidden, nb_of_layers=layers).double()
    with torch.no_grad():
        blk.edge_block[2 * layers - 1].scale.uniform_(0.5, 1.5)
        blk.node_block[2 * layers - 1].scale.uniform_(0.5, 1.5)
    n, m = 50, 300
    ei = torch.randint(0, n, (2, m))
    x = torch.randn(n, hidden, dtype=torch.float64, requires_grad=True)
    e = torch.randn(m, hidden, dtype=torch.float64, requires_grad=True)
    gx, ge = torch.randn(n, hidden, dtype=torch.float64), torch.randn(m, hidden, dtype=torch.float64)
    # oracle on the original block
    p = {k: v for k, v in blk.state_dict().items()}
    p = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    xr, er = mo.graph_net_block(x, ei, e, p, "", nb_of_layers=layers)
    ((xr * gx).sum() + (er * ge).sum()).backward()
    ref = {"x": x.grad.clone(), "e": e.grad.clone(), **{k: v.grad for k, v in p.items()}}
    x.grad = e.grad = None
    # embedded evaluation
    prm = block_param_list(blk)
    assert len(prm) == 18 and tuple(prm[0].shape) == (128, 384) and tuple(prm[9].shape) == (128, 256)
    pad = 128 - hidden
    xp = torch.nn.functional.pad(x, (0, pad))
    ep = torch.nn.functional.pad(e, (0, pad))
    xo, eo = _embedded_block(xp, ei, ep, prm, hidden)
    assert float(xo[:, hidden:].abs().max() if pad else 0.0) == 0.0          # padded columns stay exactly zero
    assert float(eo[:, hidden:].abs().max() if pad else 0.0) == 0.0
    assert mo.rel_l2(xo[:, :hidden], xr) < 1e-12 and mo.rel_l2(eo[:, :hidden], er) < 1e-12
    ((xo[:, :hidden] * gx).sum() + (eo[:, :hidden] * ge).sum()).backward()
    assert mo.rel_l2(x.grad, ref["x"]) < 1e-12 and mo.rel_l2(e.grad, ref["e"]) < 1e-12
    for name, prm_ in blk.named_parameters():
        assert mo.rel_l2(prm_.grad, ref[name]) < 1e-11, name
