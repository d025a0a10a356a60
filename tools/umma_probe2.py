"""Developer probe (round-2 groundwork): UMMA M = 64 and A-operand-from-tensor-memory on sm_100a.
Prints, per case, the error against an fp64 product under each candidate row -> TMEM-lane mapping."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200 import _lib as L  # noqa: E402

L.require_device()
g = torch.Generator().manual_seed(0)
a = (torch.randn(128, 128, generator=g) * 0.5).half()
b = (torch.randn(128, 128, generator=g) * 0.5).half()
ad, bd = a.cuda(), b.cuda()
ref = a.double() @ b.double().t()          # [128 rows of A, 128 rows of B]


def run(n, ksteps, m, a_tmem):
    d = torch.full((128, n), float("nan"), device="cuda")
    L.check(L.lib().mgn_selftest_umma2(L.p(ad), L.p(bd), L.p(d), n, ksteps, m, a_tmem, L.stream()))
    torch.cuda.synchronize()
    return d.cpu().double()


def rel(x, y):
    return float((x - y).abs().max() / y.abs().max())


for name, (n, ks, m, at) in {"ss m128": (128, 8, 128, 0), "ts m128 lo-first": (128, 8, 128, 1),
                             "ts m128 hi-first": (128, 8, 128, 2), "ss m64": (128, 8, 64, 0),
                             "ts m64 lo-first": (128, 8, 64, 1), "ss m128 n64 k4": (64, 4, 128, 0)}.items():
    d = run(n, ks, m, at)
    K = 16 * ks
    want = a.double()[:, :K] @ b.double()[:n, :K].t()
    if m == 128:
        print(f"{name:20s} rel err {rel(d, want):.3e}")
    else:
        nz = [i for i in range(128) if float(d[i].abs().max()) > 0]
        cands = {"lanes 0-63": list(range(64)),
                 "lanes 0-15,32-47,64-79,96-111": [32 * (r // 16) + r % 16 for r in range(64)],
                 "lanes 0-31,64-95": [r if r < 32 else r + 32 for r in range(64)]}
        errs = {k: rel(d[v], want[:64]) for k, v in cands.items()}
        print(f"{name:20s} nonzero lanes: {nz[:4]}..{nz[-4:]} ({len(nz)}); " + ", ".join(f"{k}: {v:.2e}" for k, v in errs.items()))
