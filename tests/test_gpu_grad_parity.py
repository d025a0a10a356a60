"""GPU: gradient parity measured against what the REFERENCE's own precision setting does on this GPU.

The reference trains with `torch.set_float32_matmul_precision("high")` (graphphysics/train.py:28): its matmuls run
with TF32 operands (10-bit mantissa, as the fp16 operands of the fused kernels).  A ReLU network's gradient is
discontinuous in the pre-activations, so ANY 10-bit-operand forward differs from the exact fp64 gradient by far more
than its forward error.  This test makes that a measured number instead of an argument:

  1. the oracle port's eager torch ops are run on the B200 with TF32 matmuls (the reference's path on this GPU), one
     step and 15 steps on the reference's cylinder mesh, and its gradient rel-L2 against the fp64 CPU oracle is printed;
  2. the fused kernels' deviation from the same fp64 gradients is asserted to be <= 1.5 x that figure (plus 2e-3), and
     below the absolute bounds TOL_GRAD_1 / TOL_GRAD_15 used by the other tests.

Gradient tolerance stated in BASELINE.md: rel-L2 <= 0.06 per step and <= 0.08 after 15 steps against the exact fp64
oracle, and never worse than 1.5 x the reference's own TF32 deviation on the same inputs."""
import os

import numpy as np
import pytest
import torch

from oracle import mgn_oracle as mo
from oracle.synth import processor_inputs

pytestmark = pytest.mark.gpu
TOL_GRAD_1, TOL_GRAD_15 = 0.06, 0.08
RATIO, SLACK = 1.5, 2e-3


def _cyl(golden_dir):
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    return torch.as_tensor(z["edge_index"].astype(np.int64)), z["points"].shape[0]


def _fp64_grads(ei, x0, e0, gx, p, K):
    p64 = {k: v.double().requires_grad_(True) for k, v in p.items()}
    x = x0.double().requires_grad_(True)
    e = e0.double().requires_grad_(True)
    xk, ek = mo.processor(x, ei, e, p64, K)
    ((xk * gx.double()).sum() + 1e-4 * (ek * ek).sum()).backward()
    g = {"g_x0": x.grad, "g_e0": e.grad}
    g.update({k: v.grad for k, v in p64.items()})
    return g


def _tf32_eager_grads(ei, x0, e0, gx, p, K):
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        pd = {k: v.cuda().requires_grad_(True) for k, v in p.items()}
        x = x0.cuda().requires_grad_(True)
        e = e0.cuda().requires_grad_(True)
        xk, ek = mo.processor(x, ei.cuda(), e, pd, K)
        ((xk * gx.cuda()).sum() + 1e-4 * (ek * ek).sum()).backward()
        torch.cuda.synchronize()
        g = {"g_x0": x.grad.cpu(), "g_e0": e.grad.cpu()}
        g.update({k: v.grad.cpu() for k, v in pd.items()})
        return g
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def _kernel_grads(ei, n, x0, e0, gx, p, K):
    from graph_physics_b200 import processor as proc
    blocks = [[p[f"processor_list.{k}.{nm}"].cuda().requires_grad_(True) for nm in proc.PARAM_NAMES] for k in range(K)]
    x = x0.cuda().requires_grad_(True)
    e = e0.cuda().requires_grad_(True)
    xk, ek = proc.run_processor(x, ei.cuda(), e, blocks)
    ((xk * gx.cuda()).sum() + 1e-4 * (ek * ek).sum()).backward()
    torch.cuda.synchronize()
    g = {"g_x0": x.grad.cpu(), "g_e0": e.grad.cpu()}
    for k in range(K):
        for nm, t in zip(proc.PARAM_NAMES, blocks[k]):
            g[f"processor_list.{k}.{nm}"] = t.grad.cpu()
    return g


@pytest.mark.parametrize("K,tol", [(1, TOL_GRAD_1), (15, TOL_GRAD_15)])
def test_gradients_no_worse_than_the_references_tf32_path(golden_dir, K, tol):
    ei, n = _cyl(golden_dir)
    p = mo.init_params(K, hidden=128, seed=0)
    x0, e0, gx = processor_inputs(n, ei.shape[1], 128, seed=0)
    ref = _fp64_grads(ei, x0, e0, gx, p, K)
    tf32 = _tf32_eager_grads(ei, x0, e0, gx, p, K)
    ours = _kernel_grads(ei, n, x0, e0, gx, p, K)
    d_tf32 = {k: mo.rel_l2(tf32[k], ref[k]) for k in ref}
    d_ours = {k: mo.rel_l2(ours[k], ref[k]) for k in ref}
    w_tf32, w_ours = max(d_tf32.values()), max(d_ours.values())
    print(f"K={K}: gradient rel-L2 vs fp64 oracle -- reference's TF32 eager path on this GPU: worst {w_tf32:.3e} "
          f"(g_x0 {d_tf32['g_x0']:.3e}, g_e0 {d_tf32['g_e0']:.3e}); fused kernels: worst {w_ours:.3e} "
          f"(g_x0 {d_ours['g_x0']:.3e}, g_e0 {d_ours['g_e0']:.3e})")
    assert w_ours < tol, (w_ours, max(d_ours, key=d_ours.get))
    # tensor by tensor: no gradient may be much further from the exact one than the reference's own path puts it
    bad = {k: (d_ours[k], d_tf32[k]) for k in ref if d_ours[k] > RATIO * max(d_tf32[k], w_tf32 / 3) + SLACK}
    assert not bad, bad
    assert w_ours <= RATIO * w_tf32 + SLACK, (w_ours, w_tf32)
