"""Developer tool: times edge_bwd alone (CUDA events) on the config-2 graph."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200 import graph as G, ops  # noqa: E402
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
K = 6     # rotate over K edge-latent versions so that inputs do not sit in L2
es = [torch.randn(E, 128, device=dev) for _ in range(K)]
P = ops.p_fwd(x, packed, b[0])
e16s = []
for e in es:
    e16 = ops.e16_image(E, dev)
    ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e16_out=e16)
    e16s.append(e16)
ges = [torch.randn(E, 128, device=dev) for _ in range(K)]
g_aggr = torch.randn(N, 128, device=dev)
dz16 = ops.e16_image(E, dev)
gb = torch.zeros(ops.GRAD_BLOCK_FLOATS, device=dev)
def run(i):
    ops.edge_bwd(None, None, ges[i % K], g_aggr, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, gb,
                 dz16=dz16, e16=e16s[i % K], inplace=True)
for i in range(3):
    run(i)
torch.cuda.synchronize()
iters = 24
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
for i in range(iters):
    ev[i][0].record(); run(i); ev[i][1].record()
torch.cuda.synchronize()
ts = sorted(a.elapsed_time(b_) * 1e3 for a, b_ in ev)
print(f"edge_bwd: median {ts[len(ts)//2]:.1f} us  min {ts[0]:.1f} us  MGN_DEBUG_SKIP={os.environ.get('MGN_DEBUG_SKIP','0')}")
