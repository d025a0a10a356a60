#!/usr/bin/env python
"""In-process liveness soak: the bench workload's fwd+bwd step (32 x cylinder mesh, 15 blocks) launched back to back for
`--seconds`, alternating eager launches, CUDA-graph replays and the forward-only path, with a host-side watchdog.  Many
more launches per GPU-second than fresh bench.py processes (tools/soak.sh covers the cold-start side).

    python tools/soak.py --seconds 60 [--copies 32]
Prints one line per mode with the number of steps that completed; exits 3 from the watchdog (with the library's hang
record) if a step takes longer than --step-timeout."""
import argparse
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=60.0)
    ap.add_argument("--copies", type=int, default=32)
    ap.add_argument("--step-timeout", type=float, default=30.0)
    args = ap.parse_args()
    from graph_physics_b200 import _lib as L, graph as G, processor as proc
    from graph_physics_b200.models import EncodeProcessDecode, Graph
    L.require_device()
    dev = torch.device("cuda", 0)
    ei1, n1 = bench.load_mesh()
    ei = G.tile_batch(ei1.to(dev), n1, args.copies)
    N, E = n1 * args.copies, ei.shape[1]
    torch.manual_seed(0)
    model = EncodeProcessDecode(15, 128, 128, 128, hidden_size=128, only_processor=True).to(dev)
    params = list(model.parameters())
    x0, e0, gx = (t.to(dev) for t in bench.synth_inputs(N, E, 0))
    state = {"t": time.time(), "what": "init", "n": 0}

    def watch():
        while True:
            time.sleep(1.0)
            if time.time() - state["t"] > args.step_timeout:
                print(f"[soak] WATCHDOG: {state['what']} step {state['n']} stalled > {args.step_timeout}s | "
                      f"{L.hang_report()}", flush=True)
                os._exit(3)
    threading.Thread(target=watch, daemon=True).start()

    def step():
        for p in params:
            p.grad = None
        out = model(Graph(x=x0, edge_index=ei, edge_attr=e0))
        out.backward(gx)

    def fwd():
        with torch.no_grad():
            model(Graph(x=x0, edge_index=ei, edge_attr=e0))

    step(); torch.cuda.synchronize()
    cap = proc.CapturedStep(step)
    capf = proc.CapturedStep(fwd)
    modes = [("eager", step, 8), ("graph", cap, 8), ("fwd_graph", capf, 16), ("eager_nosync", step, 32)]
    counts = {m[0]: 0 for m in modes}
    t_end = time.time() + args.seconds
    ref = None
    while time.time() < t_end:
        for name, fn, burst in modes:
            state.update(what=name, t=time.time())
            for i in range(burst):
                fn()
                if name != "eager_nosync":
                    torch.cuda.synchronize()
                state.update(n=counts[name] + i, t=time.time())
            torch.cuda.synchronize()
            counts[name] += burst
            state["t"] = time.time()
        g = params[0].grad
        if g is not None:
            s = float(g.double().abs().sum())
            if ref is None:
                ref = s
            elif abs(s - ref) > 5e-2 * abs(ref) or s != s:
                print(f"[soak] gradient checksum drifted: {s} vs {ref}", flush=True)
                sys.exit(4)
    print("[soak] completed steps:", counts, "no stall", flush=True)


if __name__ == "__main__":
    main()
