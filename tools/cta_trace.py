"""Developer tool (needs a -DMGN_CTA_TRACE build, MGN_LIB_PATH): per-CTA %globaltimer stamps of the chained pair
node_bwd -> node_wgrad1 -- when do the consumer's CTAs start, take their first / second tile and finish, relative to the
producer's CTAs?"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200 import _lib as L, ops  # noqa: E402

dev = "cuda"
N = 1923 * 32
torch.manual_seed(0)
ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
nb = [torch.randn(128, device=dev) * 0.05 for _ in range(4)]
nsc = torch.ones(128, device=dev)
packed = ops.pack_block(ew, nw)
x, aggr, gx = (torch.randn(N, 128, device=dev) for _ in range(3))
saves = ops.node_saves(N, dev)
ops.node_fwd(x, aggr, packed, nb[0], nb[1], nb[2], nb[3], nsc, saves=saves)
grad = torch.zeros(ops.GRAD_BLOCK_FLOATS, dtype=torch.float32, device=dev)
g_x, g_ag = torch.empty_like(x), torch.empty_like(x)
dz1 = torch.empty((N, 128), dtype=torch.float16, device=dev)
cap = 2 * 148 * 4
buf = torch.zeros(cap, dtype=torch.int64, device=dev)
for chained in (False, True):
    chain = ops.chain_state(N, dev) if chained else None

    def pair():
        ops.node_bwd(gx, saves, packed, nb[3], nsc, grad, g_x=g_x, g_aggr=g_ag, dz1=dz1, chain=chain)
        ops.node_wgrad1(dz1, x, aggr, grad, chain=chain)

    for _ in range(3):
        pair()
    torch.cuda.synchronize()
    buf.zero_()
    L.check(L.lib().mgn_debug_set_timeline(ctypes.c_void_p(buf.data_ptr()), ctypes.c_int(cap)))
    pair()
    torch.cuda.synchronize()
    L.check(L.lib().mgn_debug_set_timeline(ctypes.c_void_p(0), ctypes.c_int(0)))
    t = buf.cpu().numpy().reshape(2, 148, 4).astype(np.float64)
    t0 = t[0, :, 0].min()
    us = (t - t0) / 1e3
    pr, co = us[0], us[1]
    q = lambda a: f"min {a.min():6.1f}  p25 {np.percentile(a, 25):6.1f}  median {np.median(a):6.1f}  p75 {np.percentile(a, 75):6.1f}  max {a.max():6.1f}"
    print(f"== chained={chained} (us after the first node_bwd CTA entered its loop)")
    print("node_bwd    loop start :", q(pr[:, 0]))
    print("node_bwd    loop end   :", q(pr[:, 2]))
    print("node_bwd    exit       :", q(pr[:, 3]))
    print("node_wgrad1 start      :", q(co[:, 0]))
    print("node_wgrad1 2nd tile   :", q(co[:, 1][co[:, 1] > 0]) if (co[:, 1] > 0).any() else "-")
    print("node_wgrad1 loop end   :", q(co[:, 2]))
    print("node_wgrad1 exit       :", q(co[:, 3]))
