#!/usr/bin/env python
"""bench.py -- MGN processor edges/s, forward + backward, H=128, 15 message-passing steps
(BASELINE.json metric) on N GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 2|3|4|5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

--config picks the BASELINE.json workload (default 2 = the configuration the metric is quoted on):
  1  the WHOLE training step of the cylinder.json model (H=128, 15 steps) on one CylinderFlow test mesh: Simulator,
     encoders, processor, decoder, masked L2 loss, backward, AdamW -- replayed from one CUDA graph (1 GPU)
  2  32 x CylinderFlow mesh per GPU, data-parallel (weak scaling); the JSON line also carries a `partitioned` record:
     the config-5 family point for this N (ONE Kuhn tetrahedral grid cut over the N GPUs, halo exchange every step;
     ~0.5 M nodes / 6.9 M edges per GPU, so N = 1 is the weak-scaling base of the 159^3 mesh at N = 8)
  3  16 x DeformingPlate-shaped tet mesh with mesh + world edges per GPU (synth.config3_batch), data-parallel
  4  47^3 tet mesh, 2-hop adjacency, MeshMask keep-ratio 0.6 sub-graph (synth.config4_graph), one replica per GPU
  5  the partitioned record alone as the headline value

Workload (BASELINE.json configs[1]): per GPU a block-diagonal batch of 32 copies of
the reference's own CylinderFlow test mesh (tests/mock_vtu/cylinder_0.vtu: 1,923 nodes, 11,070
directed edges; committed as tests/golden/cylinder_mesh.npz) -> N=61,536, E=354,240; processor only
(`EncodeProcessDecode(only_processor=True)`), default nn.Linear init, x0/e0 ~ RMSNorm(N(0,1)),
upstream gradient N(0,1)/sqrt(N*H).  One step = one forward + backward through the 15 blocks.
Data-parallel over meshes for N > 1 (weak scaling: 32 meshes per GPU) with an NCCL all-reduce of
each block's gradient as soon as that block's backward is done.

One JSON line is printed by rank 0 (contract in the task description): `value` = device-resident
edges/s over all GPUs; `e2e` = same through the public module API with host inputs (pinned H2D of
x0/e0/upstream gradient every step, D2H of the scalar loss); `roofline` for the dominant kernel
(edge_bwd) from CUDA events recorded on the launch stream inside the timed region; `cpu_baseline`
= the CPU oracle (port of the reference's PyTorch path) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, STEPS_MP = 128, 15
METRIC = "mgn_processor_edges_per_s_fwd_bwd_h128_15steps"
BYTES_EDGE_STEP, BYTES_NODE_STEP = 2568, 7168          # SURVEY.md 8(d), fwd+bwd per step
FLOPS_EDGE_STEP, FLOPS_NODE_STEP = 589824, 491520
EDGE_BWD_BYTES_EDGE, EDGE_BWD_BYTES_NODE = 1540, 2048  # the dominant kernel's share


def kernel_source_hash():
    """sha256 over csrc/*.cu, *.cuh: identifies the build an ncu capture belongs to."""
    import hashlib
    d = os.path.join(ROOT, "graph_physics_b200", "csrc")
    h = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh")):
            with open(os.path.join(d, f), "rb") as fh:
                h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def measured_traffic(kernel, config):
    """`roofline.traffic`: dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed
    `ncu --set full` capture (profiles/traffic.json, written by tools/ncu_traffic.py) -- only if that capture was made
    on THIS build of the kernels (source hash) and this workload; otherwise None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        if t.get("source_hash") != kernel_source_hash():
            return None, f"profiles/traffic.json is from build {t.get('source_hash')}, not this one"
        rec = t["kernels"][kernel][f"config{config}"]
        return int(rec["dram_bytes_per_launch"]), rec.get("source", "profiles/traffic.json")
    except Exception as exc:
        return None, f"no capture for this build ({type(exc).__name__})"


def load_mesh():
    z = np.load(os.path.join(ROOT, "tests", "golden", "cylinder_mesh.npz"))
    return torch.as_tensor(z["edge_index"].astype(np.int64)), int(z["points"].shape[0])


def synth_inputs(n, e, seed):
    """x0, e0 ~ RMSNorm(N(0,1)) (what the encoders emit), upstream gradient N(0,1)/sqrt(N*H)."""
    g = torch.Generator().manual_seed(1000 + seed)

    def rms(t):
        return t / (t.norm(dim=-1, keepdim=True) * H ** -0.5 + 1e-8)
    x0 = rms(torch.randn(n, H, generator=g))
    e0 = rms(torch.randn(e, H, generator=g))
    gx = torch.randn(n, H, generator=g) / (n * H) ** 0.5
    return x0, e0, gx


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        return float(m["hbm_gbs"]), float(m.get("bf16_tflops_sustained", m["bf16_tflops"])), "measured"
    except Exception:
        return 6650.0, 1590.0, "fallback"


class Progress:
    """Flushed stderr marker before every phase, and a watchdog thread: a phase that exceeds `limit` seconds (a kernel
    that never finishes holds the process inside a CUDA call, which no signal interrupts) gets the phase name, the
    library's hang record and whatever result line exists printed, and the process exits 3 instead of sitting on the
    GPU until the caller's limit."""

    def __init__(self, limit, rank=0):
        self.limit, self.rank = float(limit), rank
        self.t00 = self.t0 = time.time()
        self.name = "start"
        self.partial = None          # result line so far (printed by the watchdog if a later phase stalls)
        self.done = False
        if self.limit > 0:
            threading.Thread(target=self._watch, daemon=True).start()

    def phase(self, name):
        self.name, self.t0 = name, time.time()
        if self.rank == 0:
            print(f"[bench +{self.t0 - self.t00:6.1f}s] {name}", file=sys.stderr, flush=True)

    def finish(self):
        self.done = True
        self.phase("done")

    def _watch(self):
        while not self.done:
            time.sleep(1.0)
            if self.done or time.time() - self.t0 <= self.limit:
                continue
            msg = f"phase '{self.name}' exceeded {self.limit:.0f} s"
            try:
                from graph_physics_b200 import _lib as L
                hr = L.hang_report()
                if hr:
                    msg += " | " + hr
            except Exception:
                pass
            print(f"[bench] WATCHDOG rank {self.rank}: {msg}", file=sys.stderr, flush=True)
            if self.rank == 0:
                line = dict(self.partial) if self.partial else {"metric": METRIC, "value": None}
                line["error"] = msg
                print(json.dumps(line), flush=True)
            os._exit(3)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for nm, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------ workloads (BASELINE.json configs)
def workload_cpu_graph(config, copies):
    """(edge_index int64 [2,E] on the CPU, N, description) of the workload's graph WITHOUT a GPU: what the reference
    arm and the cpu_baseline leg run on.  Configs 3 / 4 use the oracle's numpy graph functions (same results as the
    device set-up path, tests/test_gpu_setup.py)."""
    if config in (2, 5):
        ei1, n1 = load_mesh()
        offs = (torch.arange(copies) * n1).view(1, copies, 1)
        return (ei1.view(2, 1, -1) + offs).reshape(2, -1), n1 * copies, f"{copies} x cylinder mesh"
    from graph_physics_b200 import synth
    from oracle import graph_oracle as go
    if config == 3:
        eis, off = [], 0
        for s_ in range(copies):
            pos, cells, nt = synth.plate_sample(seed=s_, gap=0.012 + 0.014 * ((s_ * 7) % 16) / 15)
            n = pos.shape[0]
            ei = go.world_edges(go.face_to_edge(go.cells_to_faces(cells.numpy()), n), pos.numpy(), nt.numpy(), n,
                                radius=0.03)
            eis.append(torch.as_tensor(ei.astype(np.int64)) + off)
            off += n
        return torch.cat(eis, dim=1), off, f"{copies} x plate + obstacle sample (mesh + world edges)"
    if config == 4:
        g = copies          # grid edge length of the CPU sample
        ei1 = synth.kuhn_grid_edges(g, g, g)
        eik = go.k_hop_edge_index(ei1.numpy(), 2, g ** 3)
        sel = torch.randperm(g ** 3, generator=torch.Generator().manual_seed(0))[:int(0.6 * g ** 3)].numpy()
        eim = go.filter_edges(eik, sel)[0]
        return torch.as_tensor(eim.astype(np.int64)), int(sel.size), f"{g}^3 tet grid, 2-hop, MeshMask keep 0.6"
    raise ValueError(config)


def workload_device_graph(config, copies, dev):
    """(edge_index [2,E] on `dev`, N, config description dict) built with the device set-up kernels."""
    from graph_physics_b200 import graph as G, synth
    if config == 2:
        ei1, n1 = load_mesh()
        return G.tile_batch(ei1.to(dev), n1, copies), n1 * copies, {
            "workload": f"BASELINE configs[1]: {copies} x CylinderFlow mesh (1923 nodes, 11070 directed edges) per GPU, "
                        "block-diagonal batch; MGN processor (only_processor=True), H=128, 15 steps, fwd+bwd"}
    if config == 3:
        d = synth.config3_batch(copies=copies, device=dev)
        return d["edge_index"], d["num_nodes"], {
            "workload": f"BASELINE configs[2]: {copies} x DeformingPlate-shaped tet sample (plate 18x18x3 + obstacle "
                        f"7x7x6 nodes) per GPU, mesh edges {d['mesh_edges']} + world edges {d['world_edges']} "
                        "(radius 0.03), 8-dim raw edge features; MGN processor, H=128, 15 steps, fwd+bwd"}
    if config == 4:
        d = synth.config4_graph(grid=47, device=dev)
        ei, n = d["masked"]
        return ei, n, {
            "workload": f"BASELINE configs[3]: 47^3 Kuhn tet mesh (103,823 nodes, {d['one_hop_edges']} 1-hop edges), 2-hop "
                        f"adjacency ({d['full'][0].shape[1]} edges), MeshMask keep-ratio 0.6 sub-graph; one replica per "
                        "GPU; MGN processor, H=128, 15 steps, fwd+bwd"}
    raise ValueError(config)


DEFAULT_COPIES = {2: 32, 3: 16, 4: 1, 5: 1}


# ------------------------------------------------------------------ CPU oracle leg
def cpu_oracle_run(ei, n, reps, warmup, threads, budget_s=None):
    """fwd+bwd through the 15 blocks with the CPU oracle (oracle/mgn_oracle.py: the same ATen ops as the reference's
    PyTorch path) on the graph `ei`.  Stops early once `budget_s` seconds of timed passes are in (at least one).
    Returns (edges/s, seconds per pass, passes timed)."""
    from oracle import mgn_oracle as mo
    torch.set_num_threads(threads)
    e = ei.shape[1]
    p = mo.init_params(STEPS_MP, hidden=H, seed=0)
    for v in p.values():
        v.requires_grad_(True)
    x0, e0, gx = synth_inputs(n, e, 0)
    ts = []
    for it in range(warmup + reps):
        for v in p.values():
            v.grad = None
        xi, ei_ = x0.clone().requires_grad_(True), e0.clone().requires_grad_(True)
        t0 = time.perf_counter()
        x, _ = mo.processor(xi, ei, ei_, p, STEPS_MP)
        x.backward(gx)
        dt = time.perf_counter() - t0
        if it >= warmup:
            ts.append(dt)
            if budget_s is not None and sum(ts) > budget_s:
                break
    t = float(np.median(ts))
    return e / t, t, len(ts)


def eager_gpu_run(ei, n, x0, e0, gx, reps=3, warmup=1):
    """Part of the baseline leg: the oracle port's eager PyTorch ops on the GPU with TF32 enabled
    (graphphysics/train.py:28) -- what the reference's code path costs on this B200.  Returns ms per fwd+bwd."""
    from oracle import mgn_oracle as mo
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        p = {k: v.to(x0.device).requires_grad_(True) for k, v in mo.init_params(STEPS_MP, hidden=H, seed=0).items()}
        ts = []
        for it in range(warmup + reps):
            for v in p.values():
                v.grad = None
            xi, ei_ = x0.clone().requires_grad_(True), e0.clone().requires_grad_(True)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            x, _ = mo.processor(xi, ei, ei_, p, STEPS_MP)
            x.backward(gx)
            e.record()
            torch.cuda.synchronize()
            if it >= warmup:
                ts.append(s.elapsed_time(e))
        return float(np.median(ts))
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def cpu_sample_copies(args):
    """How much of the workload the CPU legs run: config 2 / 3 the whole per-GPU batch (same config as the GPU arm),
    config 4 a 16^3 grid of the same family (the 47^3 graph takes minutes per pass on the host), config 5 the
    config-2 batch (the reference cannot hold the 4 M-node mesh's autograd state on the host)."""
    if args.cpu_copies > 0:
        return args.cpu_copies
    return {2: 32, 3: 16, 4: 16, 5: 32}[args.config]


# ------------------------------------------------------------------ reference arm (CPU)
def run_reference(args):
    """`--impl reference`: the reference's own CPU path for this metric.  The reference cannot be imported here or on
    the GPU box (torch_geometric absent, no network), so the timed code is the oracle port of its PyTorch
    ops, with all host threads, on the SAME per-GPU workload as the B200 arm (all 32 meshes for config 2).  One step
    = one fwd+bwd pass (several seconds on the host): at most `--steps` passes are timed and the run stops once
    ~90 s of timed passes are in.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    prog = Progress(args.phase_timeout if args.phase_timeout > 0 else 0, 0)
    prog.limit = max(prog.limit, 900.0)
    threads = os.cpu_count() or 1
    copies = cpu_sample_copies(args)
    prog.phase("reference arm: building the workload graph on the host")
    ei, n, what = workload_cpu_graph(args.config, copies)
    prog.phase(f"reference arm: oracle port fwd+bwd on {threads} threads, E={ei.shape[1]}")
    eps, t, done = cpu_oracle_run(ei, n, max(1, args.steps), min(args.warmup, 1), threads, budget_s=90.0)
    sample = (f"{what}: N={n}, E={ei.shape[1]} per pass, {done} timed passes (median), oracle port of the reference's "
              "PyTorch ops, fp32")
    cfg = {"workload": f"BASELINE config {args.config}: {what}; MGN processor, H=128, 15 steps, fwd+bwd",
           "nodes_per_gpu": n, "edges_per_gpu": int(ei.shape[1]), "hidden": H, "mp_steps": STEPS_MP,
           "parallelism": f"dp{args.gpus}", "sample": sample}
    line = {
        "impl": "reference", "metric": METRIC, "value": eps, "unit": "edges/s", "n_gpus": args.gpus,
        "steps": done, "warmup": min(args.warmup, 1), "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": eps, "unit": "edges/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": eps, "unit": "edges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    prog.finish()


def workload_config(n_gpus, n, e, desc):
    cfg = dict(desc)
    cfg.update({
        "nodes_per_gpu": n, "edges_per_gpu": e, "hidden": H, "mp_steps": STEPS_MP, "parallelism": f"dp{n_gpus}",
        "numerics": "fp16 tensor-core operands, fp32 accumulation; residual adds, RMSNorm, aggregation and gradients in "
                    "fp32; edge stream held between steps as fp16 image + fp16 remainder (>= 21 bits)",
        "l2": f"inputs larger than L2: every step streams a {e * 512 / 1e6:.0f} MB edge-latent version and a "
              f"{e * 512 / 1e6:.0f} MB edge-gradient version (L2 126 MB)"})
    return cfg


# ------------------------------------------------------------------ config 5: one mesh cut over the GPUs
def run_partitioned_point(args, prog, dev, rank, world, grid=0, steps=5, warmup=3):
    """The config-5 family point for this world size (tools/bench_partitioned.py is the stand-alone form): ONE g^3 Kuhn
    tetrahedral grid, node-partitioned by receiver over `world` GPUs (graph_physics_b200.partition), a halo exchange
    of boundary node latents every message-passing step forward and of their gradients backward.  Default g: the cube
    closest to world * 159^3 / 8 nodes (79, 100, 126, 159 at 1, 2, 4, 8 GPUs), i.e. ~0.5 M nodes / 6.9 M edges per
    GPU -- weak scaling towards BASELINE's 4 M-node mesh with the 1-GPU point as base."""
    import torch.distributed as dist

    from graph_physics_b200 import dist_processor as DP
    from graph_physics_b200 import graph as G
    from graph_physics_b200 import partition as PT
    from graph_physics_b200 import processor as proc
    from graph_physics_b200 import synth
    from graph_physics_b200.models import EncodeProcessDecode
    from graph_physics_b200.models.layers import block_param_list

    prog.phase("partitioned: grid edges, partition, local CSR")
    g = grid or synth.grid_for_nodes(world * 159 ** 3 / 8)
    n = g ** 3
    ei = synth.kuhn_grid_edges(g, g, g, device=dev)
    E = int(ei.shape[1])
    part = PT.build_partition(ei, n, rank, world)
    del ei
    torch.manual_seed(0)
    model = EncodeProcessDecode(STEPS_MP, H, H, H, hidden_size=H, only_processor=True).to(dev)
    blocks = [block_param_list(b) for b in model.processor_list]

    def rows(ids, seed):       # functions of the GLOBAL node / edge id: every world size sees the same problem
        gen = torch.Generator(device=dev).manual_seed(seed)
        base = torch.randn((1024, H), generator=gen, device=dev)
        return base[ids % 1024] + 0.5 * base[(ids // 1024) % 1024].roll(1, dims=1)

    def rms(t):
        return t / (t.norm(dim=-1, keepdim=True) * H ** -0.5 + 1e-8)
    own_ids = torch.arange(part.node_begin, part.node_begin + part.num_owned, device=dev)
    x_own = rms(rows(own_ids, 1)).requires_grad_(True)
    e_loc = rms(rows(part.edge_ids, 2)).requires_grad_(True)
    gx = rows(own_ids, 3) / (n * H) ** 0.5
    local_csr = G.build_csr(part.edge_index, part.num_ext)
    if part.edge_index.shape[1]:
        local_csr.by_sender

    def step():
        for p_ in model.parameters():
            p_.grad = None
        x_own.grad = None
        e_loc.grad = None
        xo, _ = DP.run_partitioned(x_own, e_loc, part, blocks, csr=local_csr, need_edges=False)
        xo.backward(gx)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    prog.phase(f"partitioned: {g}^3 grid over {world} GPU(s), warm-up + timed steps")
    for _ in range(warmup):
        step()
    sync()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = proc.LAUNCHES
    t0.record()
    for _ in range(steps):
        step()
    t1.record()
    sync()
    launches = proc.LAUNCHES - l0
    ms = torch.tensor([t0.elapsed_time(t1) / steps], device=dev)
    prog.phase("partitioned: halo exchanges alone")
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
    sync()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for _ in range(STEPS_MP):
        halo_pair()
    h1.record()
    sync()
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
    t = float(ms)
    hbm, tf, which = peaks()
    step_bytes = STEPS_MP * (E * BYTES_EDGE_STEP + n * BYTES_NODE_STEP)
    rec = {
        "value": E / (t * 1e-3), "unit": "edges/s", "ms_per_step": t, "steps": steps, "warmup": warmup,
        "scaling": "weak", "n_gpus": world,
        "workload": f"BASELINE configs[4] family: {g}^3 Kuhn tetrahedral grid, N={n}, E={E}, ONE mesh node-partitioned "
                    f"by receiver over {world} GPU(s); 15 steps fwd+bwd; eager launches; 15 forward + 15 reverse halo "
                    "exchanges per step (NCCL all-to-all-v of the fp16 sender halves of P / dP: 256 B per halo node)",
        "nodes": n, "edges": E,
        "per_rank": {"owned_nodes": [int(s_[0]) for s_ in allst], "halo_nodes": [int(s_[1]) for s_ in allst],
                     "edges": [int(s_[2]) for s_ in allst], "max_mem_gib": [round(float(s_[3]), 1) for s_ in allst]},
        "halo_exchange_ms_per_step": float(halo_ms), "halo_exchanges_per_step": 2 * STEPS_MP,
        "gpu_launches": launches,
        "step_hbm_frac_per_gpu": step_bytes / world / (t * 1e-3) / 1e9 / hbm,
    }
    del x_own, e_loc, gx, local_csr, part, model, blocks, pbuf, recv_f, recv_b
    torch.cuda.empty_cache()
    return rec


# ------------------------------------------------------------------ config 1: the whole training step
def config1_frames(copies, dev=None):
    """CylinderFlow-shaped raw frames on `copies` block-diagonal copies of the reference's test mesh (cylinder.json
    layout: x = [velocity (2) | node type], target = next velocity; node types by position as SURVEY.md 8(d): inflow at
    the left boundary, outflow at the right, walls at y extremes, the rest normal)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "cylinder_mesh.npz"))
    pos1 = torch.as_tensor(z["points"][:, :2]).float()
    ei1 = torch.as_tensor(z["edge_index"].astype(np.int64))
    n1 = pos1.shape[0]
    g = torch.Generator().manual_seed(0)
    nt1 = torch.zeros(n1)
    xmin, xmax, ymin, ymax = pos1[:, 0].min(), pos1[:, 0].max(), pos1[:, 1].min(), pos1[:, 1].max()
    nt1[(pos1[:, 1] <= ymin + 1e-6) | (pos1[:, 1] >= ymax - 1e-6)] = 6.0      # WALL_BOUNDARY
    nt1[pos1[:, 0] <= xmin + 1e-6] = 4.0                                      # INFLOW
    nt1[pos1[:, 0] >= xmax - 1e-6] = 5.0                                      # OUTFLOW
    vel = torch.randn(n1 * copies, 2, generator=g)
    x = torch.cat([vel, nt1.repeat(copies).unsqueeze(1)], dim=1)
    y = vel + 0.01 * torch.randn(n1 * copies, 2, generator=g)
    offs = (torch.arange(copies) * n1).view(1, copies, 1)
    ei = (ei1.view(2, 1, -1) + offs).reshape(2, -1).contiguous()
    pos = pos1.repeat(copies, 1)
    return x, y, pos, ei


def config1_reference_step(x, y, pos, ei, steps, warmup, threads, budget_s=90.0):
    """CPU restatement of one training step of the reference for config 1 with the oracle's ops: normalise (batch
    statistics, as the first accumulation of the reference's Normalizer), one-hot node type, Cartesian / Distance edge
    features, EncodeProcessDecode(15, 11, 3, 2) forward, masked MSE, backward, AdamW step.  Returns (edges/s, s/step)."""
    from oracle import mgn_oracle as mo
    torch.set_num_threads(threads)
    p = mo.init_params(STEPS_MP, hidden=H, node_in=11, edge_in=3, out_size=2, seed=0)
    prm = [v.requires_grad_(True) for v in p.values()]
    opt = torch.optim.AdamW(prm, lr=1e-4)
    rel = pos[ei[0]] - pos[ei[1]]
    ea = torch.cat([rel, rel.norm(dim=1, keepdim=True)], dim=1)

    def norm(t):
        return (t - t.mean(0, keepdim=True)) / t.std(0, keepdim=True, unbiased=False).clamp_min(1e-8)
    ts = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        nt = x[:, 2].long()
        feat = norm(torch.cat([x[:, :2], torch.nn.functional.one_hot(nt, 9).float()], dim=1))
        target = norm(y - x[:, :2])
        out = mo.encode_process_decode(feat, ei, norm(ea), p, STEPS_MP)
        mask = ((nt == 0) | (nt == 5)).float().unsqueeze(1)
        loss = (((out - target) ** 2) * mask).sum() / (mask.sum() * out.shape[1])
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        dt = time.perf_counter() - t0
        if it >= warmup:
            ts.append(dt)
            if sum(ts) > budget_s:
                break
    t = float(np.median(ts))
    return ei.shape[1] / t, t, len(ts)


def run_config1(args):
    """BASELINE.json configs[0]: MeshGraphNet with cylinder.json's dataset layout and H=128 / 15 message-passing steps
    on the reference's CylinderFlow test mesh -- ONE whole training step: Simulator (one-hot, three normalisers, target
    delta) -> fused encoders -> 15 processor blocks -> fused decoder -> masked L2 loss -> backward -> AdamW step, replayed
    from one CUDA graph.  `--copies` meshes per step (default 1: the reference's batch for this config)."""
    from graph_physics_b200 import _lib as L
    from graph_physics_b200 import graph as G, processor as proc
    from graph_physics_b200.models import EncodeProcessDecode, Graph, Simulator
    from graph_physics_b200.training import training_step
    from graph_physics_b200.utils import L2Loss, NodeType

    copies = args.copies if args.copies > 0 else 1
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        threads = os.cpu_count() or 1
        x, y, pos, ei = config1_frames(copies)
        eps, t, done = config1_reference_step(x, y, pos, ei, max(1, args.steps), min(args.warmup, 1), threads)
        sample = (f"{copies} x cylinder mesh, N={x.shape[0]}, E={ei.shape[1]}: full training step (normalise, one-hot, "
                  f"encoders, 15 blocks, decoder, masked MSE, backward, AdamW), oracle port, fp32, {done} timed steps")
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": eps, "unit": "edges/s", "n_gpus": args.gpus, "steps": done,
            "warmup": min(args.warmup, 1), "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE configs[0]: cylinder.json model (H=128, 15 steps) on the CylinderFlow test "
                                   f"mesh, {copies} mesh(es) per step, one whole training step", "sample": sample},
            "cpu_baseline": {"value": eps, "unit": "edges/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": eps, "unit": "edges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}), flush=True)
        return
    if int(os.environ.get("WORLD_SIZE", "1")) != 1 or args.gpus != 1:
        raise SystemExit("--config 1 is the reference's single-mesh training step: one GPU (replicas only)")
    prog = Progress(args.phase_timeout, 0)
    prog.phase("config 1: device, mesh, Simulator + model, optimiser")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    L.require_device()
    x_h, y_h, pos_h, ei_h = config1_frames(copies)
    x_h, y_h = x_h.pin_memory(), y_h.pin_memory()
    ei, pos = ei_h.to(dev), pos_h.to(dev)
    N, E = x_h.shape[0], int(ei.shape[1])
    ea = G.edge_features(pos, ei)
    torch.manual_seed(0)
    model = EncodeProcessDecode(STEPS_MP, 2 + int(NodeType.SIZE), 3, 2).to(dev)
    sim = Simulator(node_input_size=2 + int(NodeType.SIZE), edge_input_size=3, output_size=2, feature_index_start=0,
                    feature_index_end=2, output_index_start=0, output_index_end=2, node_type_index=2, model=model,
                    device=dev).train()
    loss_fn = L2Loss()
    masks = [NodeType.NORMAL, NodeType.OUTFLOW]
    params = [p for p in sim.parameters() if p.requires_grad]
    opt = torch.optim.AdamW(params, lr=1e-4, capturable=True, foreach=True)
    x, y = x_h.to(dev), y_h.to(dev)
    batch = Graph(x=x, y=y, pos=pos, edge_attr=ea, edge_index=ei)
    loss_dev = torch.zeros((), device=dev)
    loss_h = torch.empty((), dtype=torch.float32).pin_memory()

    def step():
        opt.zero_grad(set_to_none=True)
        loss = training_step(sim, batch, loss_fn, masks)
        loss.backward()
        opt.step()
        loss_dev.copy_(loss.detach())

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        l0 = proc.LAUNCHES
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            fn()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / steps, (proc.LAUNCHES - l0) // max(steps, 1)

    prog.phase("config 1: eager steps")
    ms_eager, launches = timed(step, args.steps, args.warmup)
    prog.phase("config 1: cuda-graph capture")
    cap = proc.CapturedStep(step)
    clocks = ClockSampler(0)
    clocks.start()
    prog.phase("config 1: timed region (frame resident)")
    ms, _ = timed(cap, args.steps, args.warmup)
    clk = clocks.stop()

    def e2e():        # frame from pinned host memory in, loss out (the copies are stream-ordered with the replay)
        x.copy_(x_h, non_blocking=True)
        y.copy_(y_h, non_blocking=True)
        cap()
        loss_h.copy_(loss_dev, non_blocking=True)
    prog.phase("config 1: timed region e2e (pinned host frame every step)")
    ms_e2e, _ = timed(e2e, max(3, min(args.steps, 10)), 2)
    hbm, tf, which = peaks()
    step_bytes = STEPS_MP * (E * BYTES_EDGE_STEP + N * BYTES_NODE_STEP)
    line = {
        "metric": METRIC, "value": E / (ms * 1e-3), "unit": "edges/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16", "data": "synthetic",
        "config": {"workload": "BASELINE configs[0]: cylinder.json model (2+9 node inputs, 3 edge inputs, 2 outputs) with "
                               f"H=128 / 15 steps on the CylinderFlow test mesh, {copies} mesh(es) per step (N={N}, "
                               f"E={E}): Simulator + encoders + processor + decoder + masked L2 loss + backward + AdamW",
                   "nodes_per_gpu": N, "edges_per_gpu": E, "hidden": H, "mp_steps": STEPS_MP, "parallelism": "dp1",
                   "l2": "this configuration fits L2 (11,070 edges = 87 tiles < 148 SMs): launch- and latency-bound by "
                         "construction; the frame is re-copied from pinned host memory every step in the e2e arm"},
        "e2e": {"value": E / (ms_e2e * 1e-3), "unit": "edges/s", "h2d_bytes_per_step": int(x_h.numel() + y_h.numel()) * 4,
                "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e},
        "gpu_launches": launches, "launch_mode": "cuda_graph_replay", "ms_per_step_eager": ms_eager, "clocks": clk,
        "roofline": {"bound": "hbm", "kernel": "whole step", "achieved": step_bytes / (ms * 1e-3) / 1e9, "peak": hbm,
                     "unit": "GB/s", "frac": step_bytes / (ms * 1e-3) / 1e9 / hbm, "traffic": None,
                     "peak_source": which, "note": "processor algorithmic bytes only (SURVEY 8(d)) over the whole step"},
        "loss": float(loss_dev), "kernel_source_hash": kernel_source_hash(),
    }
    prog.partial = line
    if args.cpu_seconds > 0:
        prog.phase("config 1: cpu_baseline leg")
        threads = os.cpu_count() or 1
        eps, t, done = config1_reference_step(x_h, y_h, pos_h, ei_h, 3, 1, threads, budget_s=30.0)
        line["cpu_baseline"] = {"value": eps, "unit": "edges/s", "cores": threads, "kind": "port",
                                "sample": f"the same training step restated with the oracle's ops on the host, {done} "
                                          f"timed steps, {t:.2f} s per step"}
    print(json.dumps(line), flush=True)
    prog.finish()


# ------------------------------------------------------------------ B200 arm
def run_b200(args):
    import torch.distributed as dist

    from graph_physics_b200 import _lib as L
    from graph_physics_b200 import parallel, processor as proc
    from graph_physics_b200.models import EncodeProcessDecode, Graph

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    prog = Progress(args.phase_timeout, rank)
    prog.phase("init: device, process group, library")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    L.require_device()

    if args.config == 5:
        rec = run_partitioned_point(args, prog, dev, rank, world, grid=args.grid, steps=max(3, min(args.steps, 10)),
                                    warmup=args.warmup)
        if rank == 0:
            line = {"metric": METRIC, "value": rec["value"], "unit": "edges/s", "n_gpus": world, "steps": rec["steps"],
                    "warmup": rec["warmup"], "ms_per_step": rec["ms_per_step"], "higher_is_better": True,
                    "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                    "config": {"workload": rec["workload"], "parallelism": f"node-partition x{world}",
                               "l2": "inputs larger than L2"},
                    "gpu_launches": rec["gpu_launches"], "launch_mode": "eager", "partitioned": rec,
                    "e2e": None, "roofline": None, "cpu_baseline": None}
            print(json.dumps(line), flush=True)
        prog.phase("teardown")
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        prog.finish()
        return

    prog.phase(f"set-up: config {args.config} graph (device set-up kernels), model, pinned inputs")
    copies = args.copies if args.copies > 0 else DEFAULT_COPIES[args.config]
    ei, N, desc = workload_device_graph(args.config, copies, dev)
    E = int(ei.shape[1])
    torch.manual_seed(0)                                   # same weights on every rank
    model = EncodeProcessDecode(STEPS_MP, H, H, H, hidden_size=H, only_processor=True).to(dev)
    params = list(model.parameters())
    x0_h, e0_h, gx_h = synth_inputs(N, E, rank)
    x0_h, e0_h, gx_h = x0_h.pin_memory(), e0_h.pin_memory(), gx_h.pin_memory()
    x0, e0, gx = x0_h.to(dev), e0_h.to(dev), gx_h.to(dev)
    # data-parallel gradient sum.  Eager launches: the per-block hook of the fused backward (asynchronous all-reduce
    # of block k as soon as its gradient is complete, overlapping the rest of the sweep).  CUDA-graph replay: NCCL
    # calls are kept out of the capture; one all-reduce over the flat gradient buffer follows every replay.
    sync = parallel.GradSync(average=True).install() if (world > 1 and args.no_cuda_graph) else None

    def after_replay():
        if world > 1 and sync is None:
            parallel.allreduce_last_grad_blocks(average=True)

    def step_resident():
        for p in params:
            p.grad = None
        out = model(Graph(x=x0, edge_index=ei, edge_attr=e0))
        out.backward(gx)
        if sync is not None:
            sync.wait()

    # ---- end-to-end arm: every step's inputs come from pinned host memory and its loss goes back to the host.
    # Two device input slots: the copy stream fills the next step's slot while the current step computes.
    loss_h = torch.empty((), dtype=torch.float32).pin_memory()

    class E2E:
        def __init__(self):
            self.copy_stream = torch.cuda.Stream()
            self.slots = [tuple(torch.empty_like(t) for t in (x0, e0, gx)) for _ in range(2)]
            self.ready = [torch.cuda.Event() for _ in range(2)]
            self.free = [torch.cuda.Event() for _ in range(2)]
            for ev_ in self.free:
                ev_.record()
            self.compute = [self._make(k) for k in range(2)]
            self.i = 0
            self._prefetch(0)

        def _make(self, k):
            xd, ed, gd = self.slots[k]

            def run():
                for p in params:
                    p.grad = None
                out = model(Graph(x=xd, edge_index=ei, edge_attr=ed))
                loss = (out * gd).sum()
                loss.backward()
                if sync is not None:
                    sync.wait()
                loss_h.copy_(loss.detach(), non_blocking=True)
            return run

        def capture(self):
            self.compute = [proc.CapturedStep(f) for f in self.compute]

        def _prefetch(self, k):
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(self.free[k])
                for dst, src in zip(self.slots[k], (x0_h, e0_h, gx_h)):
                    dst.copy_(src, non_blocking=True)
                self.ready[k].record(self.copy_stream)

        def __call__(self):
            k = self.i & 1
            self._prefetch(k ^ 1)                       # next step's inputs: H2D under this step's kernels
            cur = torch.cuda.current_stream()
            cur.wait_event(self.ready[k])
            self.compute[k]()
            after_replay()
            self.free[k].record(cur)
            self.i += 1

    step_e2e = E2E()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, with_events=False):
        for _ in range(warmup):
            fn()
        barrier()
        proc.EVENTS = {"edge_bwd": [], "edge_fwd": []} if with_events else None
        l0 = proc.LAUNCHES
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            fn()
        e.record()
        barrier()
        ms = s.elapsed_time(e)
        launches = proc.LAUNCHES - l0
        ev = proc.EVENTS
        proc.EVENTS = None
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches, ev

    # (1) eager pass: per-launch CUDA events of the dominant kernel (roofline) and the launch count
    def step_eager():
        step_resident()
        after_replay()

    prog.phase("csr build + eager pass (per-launch events of the edge kernels)")
    ms_eager, launches, ev = timed(step_eager, args.steps, args.warmup, with_events=True)
    # (2) the timed region of `value`: the same step replayed from a CUDA graph (no host launch gaps)
    step_fn, e2e_fn, graphed = step_eager, step_e2e, False
    if not args.no_cuda_graph:
        prog.phase("cuda-graph capture (resident step, e2e steps)")
        try:
            cap_step = proc.CapturedStep(step_resident)

            def step_fn():
                cap_step()
                after_replay()
            step_e2e.capture()
            graphed = True
        except Exception as exc:      # capture is an optimisation: report and fall back to eager launches
            if rank == 0:
                print(f"bench: CUDA-graph capture failed ({type(exc).__name__}: {exc}); timing eager launches",
                      file=sys.stderr)
            step_fn, e2e_fn, graphed = step_eager, step_e2e, False
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    prog.phase("timed region: value (inputs resident)")
    ms, _, _ = timed(step_fn, args.steps, args.warmup)
    clk = clocks.stop() if rank == 0 else None
    e2e_steps = max(2, min(args.steps, 5))
    prog.phase("timed region: e2e (pinned host inputs every step)")
    ms_e2e, _, _ = timed(e2e_fn, e2e_steps, 2)
    # forward only (inference / rollout step, SURVEY.md 8(d) "also report"): no_grad, ping-pong buffers, graph replay
    def fwd_only():
        with torch.no_grad():
            return model(Graph(x=x0, edge_index=ei, edge_attr=e0))
    fwd_fn = fwd_only
    prog.phase("forward-only region")
    if graphed:
        try:
            fwd_fn = proc.CapturedStep(fwd_only)
        except Exception:
            fwd_fn = fwd_only
    ms_fwd, _, _ = timed(fwd_fn, args.steps, 3)

    line = None
    if rank == 0:
        hbm, tf, which = peaks()
        ms_step = ms / args.steps
        value = world * E / (ms_step * 1e-3)
        durs = [a.elapsed_time(b) for a, b in ev["edge_bwd"]]
        edge_bwd_ms = float(np.mean(durs)) if durs else float("nan")
        dfw = [a.elapsed_time(b) for a, b in ev["edge_fwd"]]
        edge_fwd_ms = float(np.mean(dfw)) if dfw else float("nan")
        alg_bytes = E * EDGE_BWD_BYTES_EDGE + N * EDGE_BWD_BYTES_NODE
        ach = alg_bytes / (edge_bwd_ms * 1e-3) / 1e9
        step_bytes = STEPS_MP * (E * BYTES_EDGE_STEP + N * BYTES_NODE_STEP)
        step_flops = STEPS_MP * (E * FLOPS_EDGE_STEP + N * FLOPS_NODE_STEP)
        traffic, traffic_src = measured_traffic("edge_bwd_saved_kernel", args.config) if copies == DEFAULT_COPIES[args.config] \
            else (None, "non-default batch")
        line = {
            "metric": METRIC, "value": value, "unit": "edges/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16",
            "data": "synthetic", "config": workload_config(world, N, E, desc),
            "e2e": {"value": world * E / (ms_e2e / e2e_steps * 1e-3), "unit": "edges/s",
                    "h2d_bytes_per_step": int(x0_h.numel() + e0_h.numel() + gx_h.numel()) * 4,
                    "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / e2e_steps},
            "gpu_launches": launches,
            "launch_mode": "cuda_graph_replay" if graphed else "eager",
            "ms_per_step_eager": ms_eager / args.steps,
            "forward_only": {"value": world * E / (ms_fwd / args.steps * 1e-3), "unit": "edges/s",
                             "ms_per_step": ms_fwd / args.steps, "what": "15 blocks forward under no_grad"},
            "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "edge_bwd_saved_kernel", "achieved": ach, "peak": hbm, "unit": "GB/s",
                         "frac": ach / hbm, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": which,
                         "kernel_note": "the edge backward of steps 1..14 (saved activations); step 0's launch "
                                        "(edge_bwd_kernel, recompute) is in the same average",
                         "ms_per_launch": edge_bwd_ms, "launches_timed": len(durs),
                         "timed_in": "eager pass of the same steps (one CUDA event pair per launch) run before the "
                                     "graph-replayed region of `value`",
                         "algorithmic_bytes_per_launch": alg_bytes,
                         "edge_fwd_ms_per_launch": edge_fwd_ms},
            "step_roofline": {"hbm_frac": step_bytes / (ms_step * 1e-3) / 1e9 / hbm,
                              "tensor_frac": step_flops / (ms_step * 1e-3) / 1e12 / tf,
                              "algorithmic_GB_per_step": step_bytes / 1e9, "algorithmic_TFLOP_per_step": step_flops / 1e12,
                              "edge_bwd_share_of_step": edge_bwd_ms * STEPS_MP / ms_step},
            "kernel_source_hash": kernel_source_hash(),
        }
        # the GPU numbers exist from here on: a stall in the later legs cannot take them along (the watchdog prints
        # this line); stderr carries it at once
        prog.partial = line
        print("[bench] result before the partitioned / baseline legs: " + json.dumps(line), file=sys.stderr, flush=True)
        if os.environ.get("MGN_BENCH_BREAKDOWN"):       # developer aid: per-kernel device time of one extra step (stderr)
            names = ["pack", "p_fwd", "edge_fwd", "node_fwd", "node_bwd", "node_wgrad1", "edge_bwd", "edge_wgrad1", "dp_gather", "p_bwd"]
            proc.EVENTS = {n_: [] for n_ in names}
            s0, e0_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            step_resident()
            e0_.record()
            torch.cuda.synchronize()
            evs, proc.EVENTS = proc.EVENTS, None
            tot = 0.0
            for n_ in names:
                d = [a.elapsed_time(b) for a, b in evs[n_]]
                tot += sum(d)
                print(f"  {n_:12s} n={len(d):3d} mean {1e3 * sum(d) / max(len(d), 1):8.1f} us  total {sum(d):7.3f} ms",
                      file=sys.stderr)
            print(f"  kernels {tot:.3f} ms of step {s0.elapsed_time(e0_):.3f} ms", file=sys.stderr)

    # ---- the partitioned large-mesh point (config 5 family) for this world size, as a sub-record
    if args.config == 2 and not args.no_partitioned:
        x0_keep = (x0, e0, gx)
        del step_e2e, fwd_fn, step_fn
        try:
            cap_step = None
        except Exception:
            pass
        torch.cuda.empty_cache()
        try:
            rec = run_partitioned_point(args, prog, dev, rank, world, grid=args.grid)
        except Exception as exc:
            rec = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}
            if world > 1:
                raise
        if rank == 0:
            line["partitioned"] = rec
            prog.partial = line
        del x0_keep

    if rank == 0:
        if args.cpu_seconds > 0 and world == 1:      # the CPU leg runs at N=1 only (rank 0 would hold N-1 ranks at a barrier)
            prog.phase("cpu_baseline leg (oracle port on the host cores)")
            threads = os.cpu_count() or 1
            ccopies = cpu_sample_copies(args)
            ei_c, n_c, what = workload_cpu_graph(args.config, ccopies)
            eps, t, done = cpu_oracle_run(ei_c, n_c, 2, 1, threads, budget_s=30.0)
            line["cpu_baseline"] = {"value": eps, "unit": "edges/s", "cores": threads, "kind": "port",
                                    "sample": f"{what}: N={n_c}, E={ei_c.shape[1]}, 15 steps fwd+bwd, fp32 oracle port, "
                                              f"median of {done} after 1 warm-up, {t:.2f} s per pass"}
            prog.partial = line
            prog.phase("eager-GPU baseline leg (oracle port's torch ops on this GPU, TF32)")
            try:
                ms_eg = eager_gpu_run(ei, N, x0, e0, gx)
                line["cpu_baseline"]["eager_gpu"] = {
                    "value": E / (ms_eg * 1e-3), "unit": "edges/s", "ms_per_step": ms_eg,
                    "what": "the same oracle port (eager PyTorch ops, TF32 matmuls as graphphysics/train.py:28) on "
                            "this GPU, full workload, median of 3"}
            except Exception as exc:
                line["cpu_baseline"]["eager_gpu"] = {"unavailable": f"{type(exc).__name__}: {exc}"[:200]}
        print(json.dumps(line), flush=True)
    prog.phase("teardown")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    prog.finish()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=[1, 2, 3, 4, 5], help="BASELINE.json workload (see the header)")
    ap.add_argument("--copies", type=int, default=0, help="meshes per GPU (0 = the config's own batch: 32 / 16 / 1)")
    ap.add_argument("--grid", type=int, default=0, help="partitioned point: grid edge length (0 = weak-scaling default)")
    ap.add_argument("--no-partitioned", action="store_true", help="config 2: skip the partitioned sub-record")
    ap.add_argument("--cpu-copies", type=int, default=0, help="meshes (config 4: grid edge) in the CPU sample; 0 = default")
    ap.add_argument("--cpu-seconds", type=float, default=1.0, help="0 disables the cpu_baseline leg")
    ap.add_argument("--no-cuda-graph", action="store_true", help="time eager launches instead of a graph replay")
    ap.add_argument("--phase-timeout", type=float, default=240.0,
                    help="seconds one phase may take before the watchdog reports it and exits 3 (0 = off)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.config == 1:
        run_config1(args)
        return
    if args.impl == "reference":
        run_reference(args)
        return
    try:
        run_b200(args)
    except Exception:
        # a kernel that trapped on a starved barrier surfaces as a CUDA error here: say which wait it was
        try:
            from graph_physics_b200 import _lib as L
            hr = L.hang_report()
            if hr:
                print("[bench] " + hr, file=sys.stderr, flush=True)
        except Exception:
            pass
        raise


if __name__ == "__main__":
    main()
