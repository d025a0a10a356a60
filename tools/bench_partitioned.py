#!/usr/bin/env python
"""BASELINE.json config 5: ONE synthetic Kuhn tetrahedral grid, node-partitioned over N GPUs with a halo exchange of
boundary node latents every message-passing step (graph_physics_b200.dist_processor); forward + backward through 15
blocks, edges/s over the whole mesh.

    python tools/bench_partitioned.py --gpus 1 [--grid G]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/bench_partitioned.py --gpus N [--grid G] [--check]

Default grid: weak scaling towards the 159^3 (4.02 M nodes) mesh at 8 GPUs -- the cube whose node count is closest
to N * 159^3 / 8 (79, 100, 126, 159 for N = 1, 2, 4, 8), i.e. ~0.5 M nodes / ~6.9 M edges per GPU; the N = 1 line is
the 1/8-size sub-mesh SURVEY.md 8(e) names as the scaling base.  `--check` (small grids) compares every rank's owned
rows and the summed weight gradients with the un-partitioned processor run on the same GPU.
Timing: CUDA events around K steps after W warm-ups, barrier on both sides, max over ranks.  One JSON line (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, STEPS_MP = 128, 15


def rms(t):
    return t / (t.norm(dim=-1, keepdim=True) * (t.shape[-1] ** -0.5) + 1e-8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--grid", type=int, default=0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--mp-steps", type=int, default=STEPS_MP)
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    from graph_physics_b200 import dist_processor as DP
    from graph_physics_b200 import partition as PT
    from graph_physics_b200 import processor as proc
    from graph_physics_b200 import synth
    from graph_physics_b200.models import EncodeProcessDecode
    from graph_physics_b200.models.layers import block_param_list

    g = args.grid or synth.grid_for_nodes(world * 159 ** 3 / 8)
    n = g ** 3
    ei = synth.kuhn_grid_edges(g, g, g, device=dev)
    E = ei.shape[1]
    part = PT.build_partition(ei, n, rank, world)
    csr = None
    torch.manual_seed(0)
    model = EncodeProcessDecode(args.mp_steps, H, H, H, hidden_size=H, only_processor=True).to(dev)
    blocks = [block_param_list(b) for b in model.processor_list]
    # inputs are functions of the GLOBAL node / edge id, so that every world size sees the same problem
    def rows(ids, seed):
        gen = torch.Generator(device=dev).manual_seed(seed)
        base = torch.randn((1024, H), generator=gen, device=dev)
        a, b = ids % 1024, (ids // 1024) % 1024
        return base[a] + 0.5 * base[b].roll(1, dims=1)
    own_ids = torch.arange(part.node_begin, part.node_begin + part.num_owned, device=dev)
    x_own = rms(rows(own_ids, 1)).requires_grad_(True)
    e_loc = rms(rows(part.edge_ids, 2)).requires_grad_(True)
    gx = rows(own_ids, 3) / (n * H) ** 0.5

    check = None
    if args.check:
        xf = rms(rows(torch.arange(n, device=dev), 1)).requires_grad_(True)
        ef = rms(rows(torch.arange(E, device=dev), 2)).requires_grad_(True)
        gxf = rows(torch.arange(n, device=dev), 3) / (n * H) ** 0.5
        xo, eo = proc.run_processor(xf, ei, ef, blocks)
        xo.backward(gxf)
        gw_ref = proc.LAST_GRAD_BLOCKS.clone()
        for p in model.parameters():
            p.grad = None
        check = (xo.detach(), eo.detach(), xf.grad, ef.grad, gw_ref)
    del ei
    local_csr = None

    def step():
        for p in model.parameters():
            p.grad = None
        x_own.grad = None
        e_loc.grad = None
        xo, eo = DP.run_partitioned(x_own, e_loc, part, blocks, csr=local_csr, need_edges=args.check)
        xo.backward(gx)
        return xo, eo

    from graph_physics_b200 import graph as G
    local_csr = G.build_csr(part.edge_index, part.num_ext)
    local_csr.by_sender if part.edge_index.shape[1] else None

    def rel(a, b):
        return float((a - b).norm() / b.norm().clamp_min(1e-30))

    if args.check:
        xo, eo = step()
        lo, hi = part.node_begin, part.node_begin + part.num_owned
        gw = torch.stack([torch.cat([p.grad.flatten() for p in b]) for b in blocks])
        errs = {"x": rel(xo.detach(), check[0][lo:hi]), "e": rel(eo.detach(), check[1][part.edge_ids]),
                "gx": rel(x_own.grad, check[2][lo:hi]), "ge": rel(e_loc.grad, check[3][part.edge_ids]),
                "gw": rel(gw, check[4])}
        print(f"[rank {rank}] check vs whole mesh on one GPU: {errs}", flush=True)
        bad = errs["x"] > 1e-3 or errs["e"] > 1e-3 or max(errs["gx"], errs["ge"], errs["gw"]) > 5e-2
        flag = torch.tensor([1.0 if bad else 0.0], device=dev)
        if world > 1:
            dist.all_reduce(flag)
        if float(flag) > 0:
            raise SystemExit("partitioned result differs from the whole-mesh result")

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    proc.LAUNCHES = 0
    t0.record()
    for _ in range(args.steps):
        step()
    t1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([t0.elapsed_time(t1) / args.steps], device=dev)
    # halo traffic alone: 15 forward + 15 reverse exchanges on the real buffers
    hx = PT.HaloExchange(part)
    # what the engine exchanges per message-passing step: the fp16 sender halves of P (forward) and of dP (backward)
    n_own_, n_send_ = part.num_owned, int(part.send_index.numel())
    pbuf = torch.zeros((part.num_ext, 2 * H), dtype=torch.float16, device=dev)
    recv_f = torch.empty((part.num_halo, H), dtype=torch.float16, device=dev)
    recv_b = torch.empty((n_send_, H), dtype=torch.float16, device=dev)

    def halo_pair():
        hx.rows_forward(pbuf[:n_own_].index_select(0, part.send_index)[:, H:].contiguous(), recv_f)
        pbuf[n_own_:, H:] = recv_f
        hx.rows_backward(pbuf[n_own_:, H:].contiguous(), recv_b)
        if n_send_:
            pbuf[:n_own_, H:].index_add_(0, part.send_index, recv_b)

    for _ in range(2):
        halo_pair()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for _ in range(args.mp_steps):
        halo_pair()
    h1.record()
    torch.cuda.synchronize()
    halo_ms = torch.tensor([h0.elapsed_time(h1)], device=dev)
    stats = torch.tensor([part.num_owned, part.num_halo, int(part.edge_ids.numel()),
                          torch.cuda.max_memory_allocated(dev) / 2 ** 30], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(halo_ms, op=dist.ReduceOp.MAX)
        allst = [torch.zeros_like(stats) for _ in range(world)]
        dist.all_gather(allst, stats)
    else:
        allst = [stats]
    if rank == 0:
        t = float(ms)
        line = {
            "metric": "mgn_processor_edges_per_s_fwd_bwd_h128_15steps", "value": E / (t * 1e-3), "unit": "edges/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t,
            "higher_is_better": True, "scaling": "weak" if not args.grid else "fixed-grid", "dtype": "f16/f32",
            "data": "synthetic",
            "config": {"workload": f"cfg5 family: {g}^3 Kuhn tetrahedral grid, N={n}, E={E}, one mesh node-partitioned "
                                   f"over {world} GPU(s), halo exchange every MP step", "mp_steps": args.mp_steps,
                       "launch_mode": "eager", "inputs": "larger than L2"},
            "per_rank": {"owned_nodes": [int(s[0]) for s in allst], "halo_nodes": [int(s[1]) for s in allst],
                         "edges": [int(s[2]) for s in allst], "max_mem_gib": [round(float(s[3]), 1) for s in allst]},
            "halo_exchange_ms_per_step": float(halo_ms), "gpu_launches": proc.LAUNCHES,
            "checked_against_whole_mesh": bool(args.check),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
