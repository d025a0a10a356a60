"""Developer tool: per-phase cycle counts of block 0 of the edge forward kernel (Timeline stamps)."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200 import _lib as L, graph as G, ops  # noqa: E402

z = np.load(os.path.join(ROOT, "tests", "golden", "cylinder_mesh.npz"))
n1 = z["points"].shape[0]
ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, 32)
N, E = n1 * 32, ei.shape[1]
csr = G.build_csr(ei, N)
dev = "cuda"
torch.manual_seed(0)
ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
b = [torch.randn(128, device=dev) * 0.05 for _ in range(8)]
sc = torch.ones(128, device=dev)
packed = ops.pack_block(ew, nw)
x = torch.randn(N, 128, device=dev)
e = torch.randn(E, 128, device=dev)
e2, x2 = torch.empty_like(e), torch.empty_like(x)
P = ops.p_fwd(x, packed, b[0])
aggr = torch.empty(N, 128, device=dev)
which = sys.argv[1] if len(sys.argv) > 1 else "edge"
fine = int(sys.argv[2]) if len(sys.argv) > 2 else 0      # stamps per tile of a -DMGN_FINE_TL build
cap = 11 * 19 + 8 if which != "edge_bwd" else 8 * 19 + 2
if which == "edge_pp":
    cap = 6 * 21
if which == "edge_hl":
    cap = 10 * 19
if which == "edge_bwd_saved":
    cap = 11 * 19 + 2
if which == "edge_hl_save":
    cap = 10 * 19
if which == "node_bwd":
    cap = 10 * 4 + 3
if fine:
    cap = fine * 19
ge_in = torch.randn(E, 128, device=dev)
g_aggr = torch.randn(N, 128, device=dev)
ge_out = torch.empty_like(e)
dz16 = ops.e16_image(E, dev)
gblock = torch.zeros(ops.GRAD_BLOCK_FLOATS, device=dev)
e16 = ops.e16_image(E, dev)
ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e_out=e2, aggr=aggr, e16_out=e16)
lo16, lo16b = ops.e16_image(E, dev).zero_(), ops.e16_image(E, dev)
buf = torch.zeros(cap, dtype=torch.int64, device=dev)


hs = tuple(ops.e16_image(E, dev) for _ in range(3))
if which in ("edge_bwd_saved", "edge_hl_save"):
    ops.edge_fwd_hl(e16, lo16, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, hi_out=dz16, lo_out=lo16b,
                    aggr=aggr, h_saves=hs)


nsaves = ops.node_saves(N, dev) if hasattr(ops, "node_saves") else None
if which == "node_bwd":
    ops.node_fwd(x, aggr, packed, b[4], b[5], b[6], b[7], sc, x_out=x2, saves=nsaves)
g_xout = torch.randn(N, 128, device=dev)
g_x, g_ag = torch.empty_like(x), torch.empty_like(x)
dz1n = torch.empty((N, 128), dtype=torch.float16, device=dev)


def run():
    if which == "node_bwd":
        ops.node_bwd(g_xout, nsaves, packed, b[7], sc, gblock, g_x=g_x, g_aggr=g_ag, dz1=dz1n)
    elif which == "edge_bwd_saved":
        ops.edge_bwd_saved(hs, ge_in, g_aggr, csr.receivers, N, packed, b[3], sc, gblock, dz16=dz16)
    elif which == "edge_hl_save":
        ops.edge_fwd_hl(e16, lo16, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, hi_out=dz16,
                        lo_out=lo16b, aggr=aggr, h_saves=hs)
    elif which == "edge_bwd":
        ops.edge_bwd(None, None, ge_in, g_aggr, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, gblock,
                     dz16=dz16, e16=e16, inplace=True)
    elif which == "edge_pp":
        ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e_out=e2, aggr=aggr,
                     e16_in=e16, e16_next=dz16)
    elif which.startswith("edge_img"):
        ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e_out=e2, aggr=aggr,
                     e16_in=e16, e16_next=dz16)
    elif which == "edge_hl":
        ops.edge_fwd_hl(e16, lo16, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, hi_out=dz16,
                        lo_out=lo16b, aggr=aggr)
    elif which == "edge":
        ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e_out=e2, aggr=aggr)
    else:
        ops.node_fwd(x, aggr, packed, b[4], b[5], b[6], b[7], sc, x_out=x2)


for _ in range(3):
    run()
torch.cuda.synchronize()
L.check(L.lib().mgn_debug_set_timeline(ctypes.c_void_p(buf.data_ptr()), ctypes.c_int(cap)))
run()
torch.cuda.synchronize()
L.check(L.lib().mgn_debug_set_timeline(ctypes.c_void_p(0), ctypes.c_int(0)))
evs = []
for _ in range(6):
    a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); run(); b_.record()
    evs.append((a, b_))
torch.cuda.synchronize()
print("event us per launch:", sorted(round(a.elapsed_time(b_) * 1e3, 1) for a, b_ in evs))
t = buf.cpu().numpy()
t = t[t > 0]
names = ["load", "L1 mma", "L1 epi", "L2 mma", "L2 epi", "L3 mma", "L3 epi", "L4 mma", "norm+stage", "output", "next"]
per = 11
if which == "edge_bwd":
    names = ["top", "R1", "R2+R3", "R4+norm", "D4..D2", "D1+g'", "g_e", "next"]
    per = 8
if which == "edge_pp":
    names = ["O3(k-1)", "O1(k)", "O4(k-1)", "O2(k)", "O5(k-1)", "O0(k+1)"]
    per = 6
if which in ("edge_hl", "edge_hl_save"):
    names = ["L1 mma", "L1 epi", "L2+L3", "L4+norm", "images", "st+mma1", "ldP", "scan", "red", "next"]
    per = 10
if which == "edge_bwd_saved":
    names = ["readout", "upstream", "g build", "R4 wait", "norm bwd", "D4", "D3", "D2", "D1", "stage g_e", "next"]
    per = 11
if which == "node_bwd":
    names = ["g tile", "h wait", "R4+norm", "D4", "D3", "D2", "D1a", "g_x out", "D1b+g_aggr", "next"]
    per = 10
    print("entry -> first tile", t[1] - t[0], "| after last tile -> loop done", t[-2] - t[-3], "flush", t[-1] - t[-2],
          "kernel body", t[-1] - t[0])
    t = t[1:-2]
if fine:
    per = fine
    names = [f"s{i}" for i in range(fine)]
if which.endswith("_raw"):
    print("entry->loop", t[1] - t[0], "tiles", np.diff(t[1:-1]).tolist(), "loop total", t[-2] - t[1], "kernel cycles", t[-2] - t[0])
    sys.exit(0)
if which == "edge_bwd_saved" and len(t) == 11 * 19 + 2:
    print("after last tile -> loop done", t[-2] - t[-3], "flush", t[-1] - t[-2], "kernel body", t[-1] - t[0])
if which == "edge_bwd" and len(t) == 8 * 19 + 2:
    print("after last tile -> loop done", t[-2] - t[-3], "flush", t[-1] - t[-2], "kernel body", t[-1] - t[0])
ntile = len(t) // per
print("tiles seen", ntile)
for k in range(ntile):
    s = t[k * per:(k + 1) * per]
    d = np.diff(s)
    nxt = (t[(k + 1) * per] - s[-1]) if (k + 1) * per < len(t) else 0
    print(f"tile {k:2d} total {s[-1] - s[0]:6d} | " + " ".join(f"{names[i]}={d[i]}" for i in range(per - 1)) + f" | gap {nxt}")
