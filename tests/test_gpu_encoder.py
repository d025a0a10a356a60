"""GPU: the encoder MLPs of EncodeProcessDecode on the fused node kernels (graph_physics_b200.encoder) against the
same nn.Sequential evaluated by eager PyTorch in fp64 (= build_mlp of the reference, layers.py:77-113).
Tolerances: forward rel-L2 <= 1e-3 (one MLP = one "step" of the path's tolerance).  Gradients: <= 5e-3 against the
fp16-operand oracle differentiated through the SAME ReLU active set as the kernels (the tight check, as in
tests/test_gpu_bwd_step.py), <= 6e-2 against exact fp64 autograd (fp16 operand rounding flips the sign of a few
pre-activations that are ~0: 2-3 % on the first layer's gradients, the figure one processor block shows too)."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a.double().cpu() - b.double().cpu()).norm() / b.double().cpu().norm().clamp_min(1e-30))


@pytest.mark.parametrize("rows,fin", [(1923, 11), (11070, 3), (1000, 8), (130, 4), (777, 128), (5, 16)])
def test_fused_encoder_matches_eager_module(rows, fin):
    from graph_physics_b200 import encoder as ENC
    from graph_physics_b200.models.layers import build_mlp
    torch.manual_seed(rows + fin)
    mlp = build_mlp(fin, 128, 128).cuda()
    with torch.no_grad():
        mlp[7].scale.mul_(1.0 + 0.1 * torch.randn_like(mlp[7].scale))
    assert ENC.supported(mlp, fin)
    ref = copy.deepcopy(mlp).double()
    feat = torch.randn(rows, fin, device="cuda")
    g = torch.randn(rows, 128, device="cuda") / (rows * 128) ** 0.5
    f1 = feat.clone().requires_grad_(True)
    out = ENC.run_encoder(mlp, f1)
    out.backward(g)
    f2 = feat.double().requires_grad_(True)
    want = ref(f2)
    want.backward(g.double())
    assert _rel(out.detach(), want.detach()) < 1e-3
    errs = {"input": _rel(f1.grad, f2.grad)}
    for (name, p), (_, q) in zip(mlp.named_parameters(), ref.named_parameters()):
        assert p.grad is not None and p.grad.shape == p.shape, name
        errs[name] = _rel(p.grad, q.grad)
    bad = {k: v for k, v in errs.items() if not v < 6e-2}
    assert not bad, bad
    # inference mode: same forward, nothing saved
    with torch.no_grad():
        out2 = ENC.run_encoder(mlp, feat)
    assert torch.equal(out2, out.detach())


def test_fused_encoder_shapes_and_fallback_choice():
    from graph_physics_b200 import encoder as ENC
    from graph_physics_b200.models import EncodeProcessDecode, Graph
    from graph_physics_b200.models.layers import build_mlp
    assert not ENC.supported(build_mlp(200, 128, 128), 200)             # wider than one operand tile
    assert not ENC.supported(build_mlp(8, 128, 128, layer_norm=False), 8)
    assert not ENC.supported(build_mlp(8, 128, 128, nb_of_layers=3), 8)
    assert not ENC.supported(build_mlp(8, 64, 64), 8)
    with pytest.raises(ValueError):
        ENC.run_encoder(build_mlp(8, 64, 64).cuda(), torch.zeros(4, 8, device="cuda"))
    # zero rows
    mlp = build_mlp(3, 128, 128).cuda()
    out = ENC.run_encoder(mlp, torch.zeros(0, 3, device="cuda", requires_grad=True))
    assert out.shape == (0, 128)
    out.sum().backward()
    # the module switch: fused and eager encoders give the same model output
    torch.manual_seed(0)
    model = EncodeProcessDecode(2, 11, 3, 2).cuda()
    n, e = 500, 3000
    gr = Graph(x=torch.randn(n, 11, device="cuda"), edge_index=torch.randint(0, n, (2, e), device="cuda"),
               edge_attr=torch.randn(e, 3, device="cuda"))
    a = model(gr)
    model.fused_encoders = False
    b = model(gr)
    assert _rel(a.detach(), b.detach()) < 5e-3


def test_fused_encoder_gradients_through_the_same_active_set():
    from graph_physics_b200 import encoder as ENC
    from graph_physics_b200.models.layers import build_mlp
    from oracle import mgn_oracle as mo
    from graph_physics_b200 import ops
    torch.manual_seed(3)
    rows, fin = 1923, 11
    mlp = build_mlp(fin, 128, 128).cuda()
    feat = torch.randn(rows, fin, device="cuda")
    g = torch.randn(rows, 128, device="cuda") / (rows * 128) ** 0.5
    f1 = feat.clone().requires_grad_(True)
    ENC.DEBUG_KEEP_SAVES = True
    try:
        out = ENC.run_encoder(mlp, f1)
        masks = [(ops.image_rows(h, rows) > 0).double().cpu() for h in ENC.LAST_SAVES]     # saves are tile images
    finally:
        ENC.DEBUG_KEEP_SAVES, ENC.LAST_SAVES = False, None
    out.backward(g)
    p64 = {f"enc.{k}": v.detach().double().cpu().requires_grad_(True) for k, v in mlp.state_dict().items()}
    f2 = feat.double().cpu().requires_grad_(True)
    want = mo.mlp(f2, p64, "enc", emulate_fp16=True, masks=masks)
    want.backward(g.double().cpu())
    assert _rel(out.detach(), want.detach()) < 2e-4
    errs = {"input": _rel(f1.grad, f2.grad)}
    for name, prm in mlp.named_parameters():
        errs[name] = _rel(prm.grad, p64["enc." + name].grad)
    bad = {k: v for k, v in errs.items() if not v < 5e-3}
    assert not bad, bad


@pytest.mark.parametrize("rows,dout", [(1923, 2), (500, 3), (130, 128), (5, 1)])
def test_fused_decoder_matches_eager_module(rows, dout):
    """decode_module = build_mlp(128, 128, out, layer_norm=False) (processors.py:85-90) on the node kernels with the
    no-RMSNorm flag, against the same nn.Sequential in fp64."""
    from graph_physics_b200 import encoder as ENC
    from graph_physics_b200.models.layers import build_mlp
    torch.manual_seed(rows + dout)
    mlp = build_mlp(128, 128, dout, layer_norm=False).cuda()
    assert ENC.decoder_supported(mlp) and not ENC.decoder_supported(build_mlp(128, 128, 2).cuda())
    ref = copy.deepcopy(mlp).double()
    lat = torch.randn(rows, 128, device="cuda")
    g = torch.randn(rows, dout, device="cuda") / (rows * dout) ** 0.5
    l1 = lat.clone().requires_grad_(True)
    out = ENC.run_decoder(mlp, l1)
    assert out.shape == (rows, dout)
    out.backward(g)
    l2 = lat.double().requires_grad_(True)
    want = ref(l2)
    want.backward(g.double())
    assert _rel(out.detach(), want.detach()) < 2e-3
    errs = {"input": _rel(l1.grad, l2.grad)}
    for (name, p), (_, q) in zip(mlp.named_parameters(), ref.named_parameters()):
        assert p.grad is not None and p.grad.shape == p.shape, name
        errs[name] = _rel(p.grad, q.grad)
    bad = {k: v for k, v in errs.items() if not v < 6e-2}
    assert not bad, bad
    with torch.no_grad():
        out2 = ENC.run_decoder(mlp, lat)
    assert torch.equal(out2, out.detach())


def test_fused_encoder_survives_huge_normalised_features():
    """A normalised column at 1e6 (std still at the Normalizer's 1e-8 floor): the fused encoder must stay finite (inputs
    are clamped to the fp16 range before they become operands) and agree with the eager fp32 module on the clamped
    input; rows without such a value are unaffected."""
    from graph_physics_b200 import encoder as enc
    from graph_physics_b200.models.layers import build_mlp
    torch.manual_seed(0)
    mlp = build_mlp(11, 128, 128).cuda()
    feat = torch.randn(700, 11, device="cuda")
    feat[::7, 3] = 1.0e6
    feat[3::7, 5] = -3.0e7
    out = enc.run_encoder(mlp, feat)
    assert torch.isfinite(out).all()
    ref = mlp(feat.clamp(-65504.0, 65504.0))
    err = float((out - ref).norm() / ref.norm())
    assert err < 2e-3, err
    clean = torch.ones(700, dtype=torch.bool, device="cuda")
    clean[::7] = False
    clean[3::7] = False
    ref_exact = mlp(feat)
    assert float((out[clean] - ref_exact[clean]).norm() / ref_exact[clean].norm()) < 2e-3
    # and training through it stays finite
    feat2 = feat.clone().requires_grad_(True)
    enc.run_encoder(mlp, feat2).square().mean().backward()
    assert torch.isfinite(feat2.grad).all() and all(torch.isfinite(p.grad).all() for p in mlp.parameters())
