"""Developer tool: times the forward step kernels (p / edge / node) on the config-2 graph
(32 x the 1,923-node cylinder mesh) with CUDA events.  Not the contract bench (bench.py)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from graph_physics_b200 import graph as G, ops  # noqa: E402


def main():
    copies = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    z = np.load(os.path.join(ROOT, "tests", "golden", "cylinder_mesh.npz"))
    n1 = z["points"].shape[0]
    ei1 = torch.as_tensor(z["edge_index"].astype(np.int64)).cuda()
    ei = G.tile_batch(ei1, n1, copies)
    N, E = n1 * copies, ei.shape[1]
    csr = G.build_csr(ei, N)
    torch.manual_seed(0)
    dev = "cuda"
    ew = [torch.randn(128, 384, device=dev) * 0.05] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    nw = [torch.randn(128, 256, device=dev) * 0.06] + [torch.randn(128, 128, device=dev) * 0.08 for _ in range(3)]
    b = [torch.randn(128, device=dev) * 0.05 for _ in range(8)]
    sc = torch.ones(128, device=dev)
    packed = ops.pack_block(ew, nw)
    x = torch.randn(N, 128, device=dev)
    e = torch.randn(E, 128, device=dev)
    e2 = torch.empty_like(e)
    x2 = torch.empty_like(x)
    P = torch.empty(N, 256, dtype=torch.float16, device=dev)
    aggr = torch.empty(N, 128, device=dev)

    def step():
        ops.p_fwd(x, packed, b[0], out=P)
        ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e_out=e2, aggr=aggr)
        ops.node_fwd(x, aggr, packed, b[4], b[5], b[6], b[7], sc, x_out=x2)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    names = ["p_fwd", "edge_fwd", "node_fwd"]
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(iters)]
    for it in range(iters):
        evs[it][0].record()
        ops.p_fwd(x, packed, b[0], out=P)
        evs[it][1].record()
        ops.edge_fwd(e, None, P, csr.senders, csr.receivers, N, packed, b[1], b[2], b[3], sc, e_out=e2, aggr=aggr)
        evs[it][2].record()
        ops.node_fwd(x, aggr, packed, b[4], b[5], b[6], b[7], sc, x_out=x2)
        evs[it][3].record()
    torch.cuda.synchronize()
    for k, nm in enumerate(names):
        ts = sorted(evs[it][k].elapsed_time(evs[it][k + 1]) for it in range(iters))
        print(f"{nm}: median {ts[len(ts) // 2] * 1e3:.1f} us  min {ts[0] * 1e3:.1f} us")
    tot = sorted(evs[it][0].elapsed_time(evs[it][3]) for it in range(iters))
    med = tot[len(tot) // 2]
    print(f"N={N} E={E} fwd step median {med * 1e3:.1f} us -> {E / med / 1e3:.1f} M edge-steps/s fwd")
    ebytes = E * 1028
    print(f"edge_fwd algorithmic GB/s: {ebytes / (sorted(evs[it][1].elapsed_time(evs[it][2]) for it in range(iters))[iters // 2] * 1e-3) / 1e9:.0f}")


if __name__ == "__main__":
    main()
