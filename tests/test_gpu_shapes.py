"""GPU: hidden sizes and layer counts other than 128 / 4 through the drop-in modules:
`GraphNetBlock(hidden_size <= 128, nb_of_layers in 2..4)` and the reference's only shipped `epd` configuration,
training_config/cylinder.json:8-15 (hidden_size 32, 5 message-passing steps, 2 + 9 node inputs, 3 edge inputs, 2 outputs).
The blocks are embedded exactly into the 128-wide four-layer kernels (tests/test_cpu_padding.py proves the algebra); here
the kernels' results are compared with the fp64 oracle: rel-L2 <= 1e-3 per step forward (BASELINE.json), gradients
<= 0.10 against the exact oracle (tiny widths: one flipped ReLU unit is a visible fraction of the gradient)."""
import os

import numpy as np
import pytest
import torch

from oracle import mgn_oracle as mo

pytestmark = pytest.mark.gpu
TOL_STEP, TOL_GRAD = 1e-3, 0.10


def _cyl(golden_dir):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    return torch.as_tensor(z["edge_index"].astype(np.int64)), z["points"].shape[0], torch.tensor(z["points"][:, :2])


@pytest.mark.parametrize("hidden,layers", [(32, 4), (32, 2), (64, 3), (128, 3), (128, 2), (96, 4)])
def test_block_shapes_vs_oracle(golden_dir, hidden, layers):
    from graph_physics_b200.models import GraphNetBlock
    ei, n, _ = _cyl(golden_dir)
    torch.manual_seed(hidden + layers)
    blk = GraphNetBlock(hidden, nb_of_layers=layers).cuda()
    g = torch.Generator().manual_seed(3)
    ones = torch.ones(hidden, dtype=torch.float64)
    x0 = mo.rms_norm(torch.randn(n, hidden, generator=g, dtype=torch.float64), ones).float()
    e0 = mo.rms_norm(torch.randn(ei.shape[1], hidden, generator=g, dtype=torch.float64), ones).float()
    gx = torch.randn(n, hidden, generator=g)
    ge = torch.randn(ei.shape[1], hidden, generator=g) * 0.3
    xd, ed = x0.cuda().requires_grad_(True), e0.cuda().requires_grad_(True)
    x1, e1 = blk(xd, ei.cuda(), ed)
    assert tuple(x1.shape) == (n, hidden) and tuple(e1.shape) == (ei.shape[1], hidden)
    ((x1 * gx.cuda()).sum() + (e1 * ge.cuda()).sum()).backward()
    torch.cuda.synchronize()
    p = {k: v.detach().cpu().double().requires_grad_(True) for k, v in blk.state_dict().items()}
    xr0, er0 = x0.double().requires_grad_(True), e0.double().requires_grad_(True)
    xr, er = mo.graph_net_block(xr0, ei, er0, p, "", nb_of_layers=layers)
    ((xr * gx.double()).sum() + (er * ge.double()).sum()).backward()
    ex, ee = mo.rel_l2(x1.cpu(), xr), mo.rel_l2(e1.cpu(), er)
    assert ex < TOL_STEP and ee < TOL_STEP, (ex, ee)
    errs = {"g_x": mo.rel_l2(xd.grad.cpu(), xr0.grad), "g_e": mo.rel_l2(ed.grad.cpu(), er0.grad)}
    for name, prm in blk.named_parameters():
        assert prm.grad is not None and tuple(prm.grad.shape) == tuple(prm.shape), name
        errs[name] = mo.rel_l2(prm.grad.cpu(), p[name].grad)
    bad = {k: v for k, v in errs.items() if not v < TOL_GRAD}
    assert not bad, bad
    print(f"hidden {hidden}, {layers} layers: fwd rel-L2 x {ex:.2e} e {ee:.2e}; worst gradient rel-L2 {max(errs.values()):.2e}")


def test_reference_cylinder_json_config_runs(golden_dir):
    """training_config/cylinder.json:8-15 -- the reference's own epd configuration -- through the drop-in model:
    forward close to the fp64 oracle, one optimiser step moves every parameter."""
    from graph_physics_b200 import graph as G
    from graph_physics_b200.models import EncodeProcessDecode, Graph
    ei, n, pos = _cyl(golden_dir)
    torch.manual_seed(0)
    model = EncodeProcessDecode(message_passing_num=5, node_input_size=2 + 9, edge_input_size=3, output_size=2,
                                hidden_size=32).cuda()
    g = torch.Generator().manual_seed(1)
    feat = torch.randn(n, 11, generator=g)
    eid = ei.cuda()
    ea = G.edge_features(pos.cuda(), eid)
    out = model(Graph(x=feat.cuda(), edge_index=eid, edge_attr=ea))
    assert tuple(out.shape) == (n, 2)
    # oracle with the module's own parameters (hidden 32 everywhere)
    p = {k: v.detach().cpu().double() for k, v in model.state_dict().items()}
    x = mo.mlp(feat.double(), p, "nodes_encoder")
    e = mo.mlp(ea.cpu().double(), p, "edges_encoder")
    x, e = mo.processor(x, ei, e, p, 5)
    ref = mo.mlp(x, p, "decode_module", layer_norm=False)
    err = mo.rel_l2(out.detach().cpu(), ref)
    assert err < 1e-2, err
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    before = [q.detach().clone() for q in model.parameters()]
    out.square().mean().backward()
    opt.step()
    torch.cuda.synchronize()
    for (name, q), b in zip(model.named_parameters(), before):
        assert q.grad is not None and torch.isfinite(q.grad).all(), name
        assert not torch.equal(q.detach(), b), name
    print(f"cylinder.json (H=32, 5 steps): output rel-L2 vs fp64 oracle {err:.2e}")


def test_unsupported_block_raises():
    from graph_physics_b200.models import GraphNetBlock
    with pytest.raises(ValueError):
        GraphNetBlock(256).cuda()(torch.zeros(4, 256, device="cuda"), torch.zeros(2, 3, dtype=torch.long, device="cuda"),
                                  torch.zeros(3, 256, device="cuda"))
    with pytest.raises(ValueError):
        GraphNetBlock(128, nb_of_layers=5).cuda()(torch.zeros(4, 128, device="cuda"),
                                                  torch.zeros(2, 3, dtype=torch.long, device="cuda"),
                                                  torch.zeros(3, 128, device="cuda"))
