"""The encoder and decoder MLPs of EncodeProcessDecode (`nodes_encoder`, `edges_encoder`: build_mlp(in, 128, 128) = 4 Linear +
ReLU + RMSNorm, graphphysics/models/processors.py:72-84, layers.py:77-113) on the fused node kernels.

SURVEY.md 8(f) N1.  The node kernel computes  out = x + RMSNorm(MLP([x | a]))  with a 256-wide first layer.  An
encoder is the same chain without the residual and with a narrow input, so it runs as
    x = 0,  a = features zero-padded to 128 columns,  W1' = [0 | W1 zero-padded]
through `mgn_node_fwd`, and its backward through `mgn_node_bwd` + `mgn_node_wgrad1` (dW1 is the [:, 128:128+in] slice
of the 256-wide first-layer gradient).  No new kernel: three launches forward+backward instead of ~25 eager ones, hidden
activations kept in fp16.  The zero half costs redundant tensor work and the read of a zero matrix.

Numerics are those of the processor: fp16 tensor-core operands (the input features are rounded to fp16: they are
normalised, O(1) quantities), fp32 accumulation, fp32 RMSNorm; the gradient is pre-scaled by a power of two.
"""
from __future__ import annotations

from typing import Sequence

import torch

from . import _lib as L
from . import ops

H = ops.H
# tests only: when True, the fp16 hidden activations of the last training-mode forward stay reachable here, so a test
# can differentiate the oracle through the same ReLU active set as the kernels
DEBUG_KEEP_SAVES = False
LAST_SAVES = None


FP16_MAX = 65504.0


def _pad_cols(t: torch.Tensor, cols: int) -> torch.Tensor:
    out = torch.zeros((t.shape[0], cols), dtype=torch.float32, device=t.device)
    out[:, :t.shape[1]] = t
    return out


class _EncoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feat, w1, b1, w2, b2, w3, b3, w4, b4, scale):
        L.require_device(feat)
        if feat.dim() != 2 or feat.shape[1] > H:
            raise ValueError(f"fused encoder: input width must be <= {H}, got {tuple(feat.shape)}")
        if tuple(w1.shape) != (H, feat.shape[1]) or any(tuple(w.shape) != (H, H) for w in (w2, w3, w4)):
            raise ValueError("fused encoder supports build_mlp(in, 128, 128, nb_of_layers=4, layer_norm=True) only")
        M, fin = feat.shape
        dev = feat.device
        # The features become fp16 tensor-core operands.  A normalised column whose accumulated std is still at the
        # Normalizer's 1e-8 floor (a node type not seen yet, the time column early on) can reach 1e5..1e8, which the
        # non-saturating fp16 conversion would turn into inf and the MLP into NaN; the reference stays finite there
        # (fp32 + the encoder's RMSNorm).  Values are clamped to the fp16 finite range: finite results, exact for every
        # |x| <= 65504.
        f32 = feat.detach().float()
        a = _pad_cols(f32.clamp(-FP16_MAX, FP16_MAX), H)                # the "aggr" half of the node kernel's input
        ctx.clamped = (f32.abs() > FP16_MAX) if any(ctx.needs_input_grad[:1]) else None
        xz = torch.zeros((M, H), dtype=torch.float32, device=dev)
        w1p = torch.zeros((H, 2 * H), dtype=torch.float32, device=dev)
        w1p[:, H:H + fin] = w1.detach()
        zero_e = [torch.zeros((H, 3 * H), device=dev)] + [torch.zeros((H, H), device=dev) for _ in range(3)]
        packed = ops.pack_block(zero_e, [w1p, w2.detach(), w3.detach(), w4.detach()])
        need_grad = any(ctx.needs_input_grad)
        sv = ops.node_saves(M, dev) if need_grad else None
        out = torch.empty((M, H), dtype=torch.float32, device=dev)
        if M > 0:
            ops.node_fwd(xz, a, packed, b1.detach().contiguous(), b2.detach().contiguous(), b3.detach().contiguous(),
                         b4.detach().contiguous(), scale.detach().contiguous(), x_out=out, saves=sv)
        if need_grad:
            if DEBUG_KEEP_SAVES:
                global LAST_SAVES
                LAST_SAVES = sv
            ctx.a, ctx.xz, ctx.sv, ctx.packed, ctx.fin = a, xz, sv, packed, fin
            ctx.b4, ctx.scale = b4.detach().contiguous(), scale.detach().contiguous()
        return out

    @staticmethod
    def backward(ctx, g):
        a, xz, sv, packed, fin = ctx.a, ctx.xz, ctx.sv, ctx.packed, ctx.fin
        M = a.shape[0]
        dev = a.device
        grads = torch.zeros(ops.GRAD_BLOCK_FLOATS, dtype=torch.float32, device=dev)
        g_feat = None
        if M > 0:
            amax = g.detach().abs().max()
            s = torch.exp2(torch.floor(torch.log2(64.0 / amax.clamp_min(1e-30)))).clamp(2.0 ** -100, 2.0 ** 100)
            gs = (g.detach() * s).contiguous()
            _, g_a, dz1 = ops.node_bwd(gs, sv, packed, ctx.b4, ctx.scale, grads)
            ops.node_wgrad1(dz1, xz, a, grads)
            grads.mul_(1.0 / s)
            if ctx.needs_input_grad[0]:
                g_feat = g_a[:, :fin] * (1.0 / s)
                if ctx.clamped is not None:
                    g_feat = g_feat.masked_fill(ctx.clamped, 0.0)      # d clamp / dx = 0 outside the range
        elif ctx.needs_input_grad[0]:
            g_feat = torch.zeros((0, fin), dtype=torch.float32, device=dev)
        v = ops.split_grad_block(grads)
        need = ctx.needs_input_grad
        names = ["node_block.0.weight", "node_block.0.bias", "node_block.2.weight", "node_block.2.bias",
                 "node_block.4.weight", "node_block.4.bias", "node_block.6.weight", "node_block.6.bias",
                 "node_block.7.scale"]
        out = [g_feat]
        for i, nme in enumerate(names):
            if not need[1 + i]:
                out.append(None)
            elif i == 0:
                out.append(v[nme][:, H:H + fin].contiguous())
            else:
                out.append(v[nme])
        return tuple(out)


class _DecoderFn(torch.autograd.Function):
    """`decode_module` of EncodeProcessDecode (processors.py:85-90: build_mlp(128, 128, out, layer_norm=False)) on the
    same node kernels: x = 0, a = the latents, W1' = [0 | W1], the last layer zero-padded from `out` to 128 rows, the
    kernels' no-RMSNorm flag (mgn_node_fwd2 / mgn_node_bwd2); the first `out` columns of the result are the output."""

    @staticmethod
    def forward(ctx, lat, w1, b1, w2, b2, w3, b3, w4, b4):
        L.require_device(lat)
        d = w4.shape[0]
        if lat.dim() != 2 or lat.shape[1] != H or d > H or tuple(w4.shape) != (d, H):
            raise ValueError("fused decoder supports build_mlp(128, 128, out <= 128, nb_of_layers=4, layer_norm=False)")
        if any(tuple(w.shape) != (H, H) for w in (w1, w2, w3)):
            raise ValueError("fused decoder supports hidden_size=128 only")
        M, dev = lat.shape[0], lat.device
        a = lat.detach().float().contiguous()
        xz = torch.zeros((M, H), dtype=torch.float32, device=dev)
        w1p = torch.zeros((H, 2 * H), dtype=torch.float32, device=dev)
        w1p[:, H:] = w1.detach()
        w4p = torch.zeros((H, H), dtype=torch.float32, device=dev)
        w4p[:d] = w4.detach()
        b4p = torch.zeros(H, dtype=torch.float32, device=dev)
        b4p[:d] = b4.detach()
        ones = torch.ones(H, dtype=torch.float32, device=dev)
        zero_e = [torch.zeros((H, 3 * H), device=dev)] + [torch.zeros((H, H), device=dev) for _ in range(3)]
        packed = ops.pack_block(zero_e, [w1p, w2.detach(), w3.detach(), w4p])
        need_grad = any(ctx.needs_input_grad)
        sv = ops.node_saves(M, dev) if need_grad else None
        out = torch.empty((M, H), dtype=torch.float32, device=dev)
        if M > 0:
            ops.node_fwd(xz, a, packed, b1.detach().contiguous(), b2.detach().contiguous(), b3.detach().contiguous(),
                         b4p, ones, x_out=out, saves=sv, norm=False)
        if need_grad:
            ctx.a, ctx.xz, ctx.sv, ctx.packed, ctx.d, ctx.b4p, ctx.ones = a, xz, sv, packed, d, b4p, ones
        return out[:, :d].contiguous()

    @staticmethod
    def backward(ctx, g):
        a, xz, sv, packed, d = ctx.a, ctx.xz, ctx.sv, ctx.packed, ctx.d
        M, dev = a.shape[0], a.device
        grads = torch.zeros(ops.GRAD_BLOCK_FLOATS, dtype=torch.float32, device=dev)
        g_lat = None
        if M > 0:
            amax = g.detach().abs().max()
            s = torch.exp2(torch.floor(torch.log2(64.0 / amax.clamp_min(1e-30)))).clamp(2.0 ** -100, 2.0 ** 100)
            gs = torch.zeros((M, H), dtype=torch.float32, device=dev)
            gs[:, :d] = g.detach() * s
            _, g_a, dz1 = ops.node_bwd(gs, sv, packed, ctx.b4p, ctx.ones, grads, norm=False)
            ops.node_wgrad1(dz1, xz, a, grads)
            grads.mul_(1.0 / s)
            if ctx.needs_input_grad[0]:
                g_lat = g_a * (1.0 / s)
        elif ctx.needs_input_grad[0]:
            g_lat = torch.zeros((0, H), dtype=torch.float32, device=dev)
        v = ops.split_grad_block(grads)
        need = ctx.needs_input_grad
        pick = [v["node_block.0.weight"][:, H:].contiguous(), v["node_block.0.bias"], v["node_block.2.weight"],
                v["node_block.2.bias"], v["node_block.4.weight"], v["node_block.4.bias"],
                v["node_block.6.weight"][:d].contiguous(), v["node_block.6.bias"][:d].contiguous()]
        return (g_lat,) + tuple(t if need[1 + i] else None for i, t in enumerate(pick))


def decoder_supported(mlp: torch.nn.Module) -> bool:
    """True if `mlp` is build_mlp(128, 128, out <= 128, nb_of_layers=4, layer_norm=False)."""
    mods = list(mlp)
    if len(mods) != 7:
        return False
    lin = [mods[i] for i in (0, 2, 4, 6)]
    if not all(isinstance(m, torch.nn.Linear) and m.bias is not None for m in lin):
        return False
    return (all(tuple(m.weight.shape) == (H, H) for m in lin[:3]) and lin[3].weight.shape[1] == H
            and lin[3].weight.shape[0] <= H)


def run_decoder(mlp: torch.nn.Module, lat: torch.Tensor) -> torch.Tensor:
    """`mlp(lat)` for a decoder-shaped build_mlp module on the fused kernels (raises if unsupported; no eager
    fallback here -- the caller chooses)."""
    if not decoder_supported(mlp):
        raise ValueError("fused decoder: unsupported MLP shape")
    return _DecoderFn.apply(lat, mlp[0].weight, mlp[0].bias, mlp[2].weight, mlp[2].bias, mlp[4].weight, mlp[4].bias,
                            mlp[6].weight, mlp[6].bias)


def supported(mlp: torch.nn.Module, in_features: int) -> bool:
    """True if `mlp` is build_mlp(in <= 128, 128, 128, nb_of_layers=4, layer_norm=True)."""
    mods = list(mlp)
    if len(mods) != 8 or in_features > H:
        return False
    lin = [mods[i] for i in (0, 2, 4, 6)]
    if not all(isinstance(m, torch.nn.Linear) for m in lin) or not hasattr(mods[7], "scale"):
        return False
    if getattr(mods[7], "bias", False) or getattr(mods[7], "p", -1.0) >= 0.0:
        return False
    return (tuple(lin[0].weight.shape) == (H, in_features)
            and all(tuple(m.weight.shape) == (H, H) for m in lin[1:]) and all(m.bias is not None for m in lin))


def encoder_params(mlp: torch.nn.Module) -> Sequence[torch.Tensor]:
    return [mlp[0].weight, mlp[0].bias, mlp[2].weight, mlp[2].bias, mlp[4].weight, mlp[4].bias, mlp[6].weight,
            mlp[6].bias, mlp[7].scale]


def run_encoder(mlp: torch.nn.Module, feat: torch.Tensor) -> torch.Tensor:
    """`mlp(feat)` for an encoder-shaped build_mlp module on the fused kernels (raises if the shape is not supported
    or the tensors are not on an sm_100 device: no eager fallback here -- the caller chooses)."""
    if not supported(mlp, feat.shape[1]):
        raise ValueError("fused encoder: unsupported MLP shape")
    return _EncoderFn.apply(feat, *encoder_params(mlp))
