"""GPU: pins the tcgen05 shared-memory descriptor conventions of csrc/tc05.cuh
(one 128-row tile product per case) against an fp64 CPU product."""
import ctypes

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
