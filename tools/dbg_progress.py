import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, "/root/repo")
from graph_physics_b200 import _lib as L, graph as G, ops
z = np.load("/root/repo/tests/golden/cylinder_mesh.npz")
n1 = z["points"].shape[0]
copies = int(sys.argv[1]) if len(sys.argv) > 1 else 32
ei = G.tile_batch(torch.as_tensor(z["edge_index"].astype(np.int64)).cuda(), n1, copies)
N, E = n1 * copies, ei.shape[1]
csr = G.build_csr(ei, N)
dev = "cuda"
torch.manual_seed(0)
ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
b = [torch.randn(128, device=dev) * 0.05 for _ in range(8)]
sc = torch.ones(128, device=dev)
packed = ops.pack_block(ew, nw)
x = torch.randn(N, 128, device=dev); e = torch.randn(E, 128, device=dev)
P = ops.p_fwd(x, packed, b[0])
e16 = ops.e16_image(E, dev)
e2, aggr = ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e16_out=e16)
ge = torch.randn(E, 128, device=dev); g_aggr = torch.randn(N, 128, device=dev)
gblock = torch.zeros(ops.GRAD_BLOCK_FLOATS, device=dev)
torch.cuda.synchronize()
buf = torch.zeros(256, dtype=torch.int64).pin_memory()
L.check(L.lib().mgn_debug_set_timeline(ctypes.c_void_p(buf.data_ptr()), ctypes.c_int(256)))
try:
    ops.edge_bwd(None, None, ge, g_aggr, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, gblock, e16=e16, inplace=True)
    torch.cuda.synchronize()
    print("OK")
except Exception as ex:
    print("FAILED", str(ex)[:80])
t = buf.numpy(); print("stamps written:", int((t > 0).sum()), "(9 per tile)")
