"""GPU: pins the tcgen05 shared-memory descriptor conventions of csrc/tc05.cuh
(one 128-row tile product per case) against an fp64 CPU product."""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = {
    # name: (a_cols, b_cols, a_mn, b_mn, n, ksteps, b_col0, use_ones, repeat)
    "kmajor_kmajor": (128, 128, 0, 0, 128, 8, 0, 0, 1),
    "mn_mn_wgrad": (128, 128, 1, 1, 128, 8, 0, 0, 1),
    "k_mn_dgrad": (128, 128, 0, 1, 128, 8, 0, 0, 1),
    "mn_ones_bias": (128, 128, 1, 0, 16, 8, 0, 1, 1),
    "accumulate_x2": (128, 128, 0, 0, 128, 8, 0, 0, 2),
    "k256_kmajor": (256, 256, 0, 0, 128, 16, 0, 0, 1),
    "mn_n256_wgrad": (128, 256, 1, 1, 256, 8, 0, 0, 1),
    "mn_col0_128": (128, 256, 1, 1, 128, 8, 128, 0, 1),
    "k_mn_n64_cols": (128, 128, 0, 1, 64, 8, 64, 0, 1),
}


def expected(a, b, cfg):
    a_cols, b_cols, a_mn, b_mn, n, ksteps, b_col0, use_ones, repeat = cfg
    A = a.double()
    B = b.double()
    K = 16 * ksteps
    Am = A[:K, :128].t() if a_mn else A[:, :K]            # [128, K]
    if use_ones:
        Bm = torch.ones(n, K, dtype=torch.float64)
    else:
        Bm = B[:K, b_col0:b_col0 + n].t() if b_mn else B[:n, :K]  # [n, K]
    return repeat * (Am @ Bm.t())


def run_case(cfg, lbo_a=-1, sbo_a=-1, lbo_b=-1, sbo_b=-1, seed=0):
    from graph_physics_b200 import _lib as L
    L.require_device()
    a_cols, b_cols, a_mn, b_mn, n, ksteps, b_col0, use_ones, repeat = cfg
    g = torch.Generator().manual_seed(seed)
    a = (torch.randn(128, a_cols, generator=g) * 0.5).half()
    b = (torch.randn(128, b_cols, generator=g) * 0.5).half()
    d = torch.zeros(128, n, dtype=torch.float32, device="cuda")
    ad, bd = a.cuda(), b.cuda()
    L.check(L.lib().mgn_selftest_umma(L.p(ad), L.p(bd), L.p(d), a_cols, b_cols, a_mn, b_mn, n, ksteps, b_col0,
                                      use_ones, repeat, lbo_a, sbo_a, lbo_b, sbo_b, L.stream()))
    torch.cuda.synchronize()
    ref = expected(a, b, cfg)
    err = float((d.cpu().double() - ref).abs().max() / ref.abs().max())
    return err


@pytest.mark.parametrize("name", list(CASES))
def test_umma_case(name):
    err = run_case(CASES[name])
    assert err < 2e-3, f"{name}: rel max err {err}"


def _run2(n, ksteps, m, a_tmem, seed=0):
    from graph_physics_b200 import _lib as L
    L.require_device()
    g = torch.Generator().manual_seed(seed)
    a = (torch.randn(128, 128, generator=g) * 0.5).half()
    b = (torch.randn(128, 128, generator=g) * 0.5).half()
    d = torch.full((128, n), float("nan"), device="cuda")
    ad, bd = a.cuda(), b.cuda()
    L.check(L.lib().mgn_selftest_umma2(L.p(ad), L.p(bd), L.p(d), n, ksteps, m, a_tmem, L.stream()))
    torch.cuda.synchronize()
    K = 16 * ksteps
    return d.cpu().double(), a.double()[:, :K] @ b.double()[:n, :K].t()


def test_a_operand_from_tensor_memory():
    """tcgen05.mma with the A operand in TMEM (written by tcgen05.st): row r = TMEM lane r, one k-step = 8 32-bit
    columns, K element 2c in the LOW half of column c.  Not used by the fused kernels yet."""
    for n, ks in ((128, 8), (64, 4)):
        d, want = _run2(n, ks, 128, 1)
        assert float((d - want).abs().max() / want.abs().max()) < 2e-3
    d, want = _run2(128, 8, 128, 2)                     # the other packing is wrong, so the probe is discriminating
    assert float((d - want).abs().max() / want.abs().max()) > 0.1


def test_m64_accumulator_lanes():
    """UMMA M = 64 (cta_group::1): accumulator row r lands in TMEM lane 32 * (r // 16) + r % 16; the other lanes
    are not written."""
    d, want = _run2(128, 8, 64, 0)
    lanes = [32 * (r // 16) + r % 16 for r in range(64)]
    assert float((d[lanes] - want[:64]).abs().max() / want.abs().max()) < 2e-3
    rest = [i for i in range(128) if i not in lanes]
    assert float(d[rest].abs().max()) == 0.0


@pytest.mark.parametrize("n,ksteps,a_tmem", [(128, 8, 0), (128, 8, 1), (256, 8, 0), (64, 4, 1)])
def test_cta_pair_product(n, ksteps, a_tmem):
    """tcgen05.mma.cta_group::2 on a cluster of two CTAs (selftest.cu: umma_selftest_pair_kernel): M = 256, every CTA
    holds its 128 rows of A and HALF of B's rows -- the weight split DESIGN.md plans for the edge kernels."""
    from graph_physics_b200 import _lib as L
    L.require_device()
    g = torch.Generator().manual_seed(n + ksteps + a_tmem)
    a = (torch.randn(256, 128, generator=g) * 0.5).half()
    b = (torch.randn(n, 128, generator=g) * 0.5).half()
    d = torch.full((256, n), float("nan"), device="cuda")
    ad, bd = a.cuda(), b.cuda()
    L.check(L.lib().mgn_selftest_umma_pair(L.p(ad), L.p(bd), L.p(d), n, ksteps, a_tmem, L.stream()))
    torch.cuda.synchronize()
    K = 16 * ksteps
    want = a.double()[:, :K] @ b.double()[:, :K].t()
    assert float((d.cpu().double() - want).abs().max() / want.abs().max()) < 2e-3
