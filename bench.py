#!/usr/bin/env python
"""bench.py -- MGN processor edges/s, forward + backward, H=128, 15 message-passing steps
(BASELINE.json metric) on N GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

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
# dram__bytes_read.sum + dram__bytes_write.sum of one edge_bwd_kernel launch on this workload, from the committed
# `ncu --set full` capture (profiles/r01_ncu_edge_bwd_v6.md); not re-measured by bench.py
EDGE_BWD_DRAM_BYTES_NCU = 338_405_632 + 209_580_800


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


# ------------------------------------------------------------------ CPU oracle leg
def cpu_oracle_run(copies, reps, warmup, threads):
    """fwd+bwd through the 15 blocks with the CPU oracle (oracle/mgn_oracle.py: the same ATen ops as
    the reference's PyTorch path) on `copies` meshes.  Returns (edges/s, seconds per pass, E)."""
    from oracle import mgn_oracle as mo
    torch.set_num_threads(threads)
    ei1, n1 = load_mesh()
    offs = (torch.arange(copies) * n1).view(1, copies, 1)
    ei = (ei1.view(2, 1, -1) + offs).reshape(2, -1)
    n, e = n1 * copies, ei.shape[1]
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
    t = float(np.median(ts))
    return e / t, t, e


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


# ------------------------------------------------------------------ reference arm (CPU)
def run_reference(args):
    """`--impl reference`: the reference's own CPU path for this metric.  The reference cannot be
    imported here or on the GPU box (torch_geometric absent, no network), so the timed
    code is the oracle port of its PyTorch ops, with all host threads, on a bounded sample (one
    mesh of the 32-mesh batch).  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    copies = args.cpu_copies
    eps, t, e = cpu_oracle_run(copies, args.steps, args.warmup, threads)
    sample = f"{copies} of 32 meshes per step ({e} edges), oracle port of the reference's PyTorch ops, fp32"
    line = {
        "impl": "reference", "metric": METRIC, "value": eps, "unit": "edges/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus, copies_note=sample),
        "cpu_baseline": {"value": eps, "unit": "edges/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": eps, "unit": "edges/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus, copies=32, copies_note=None):
    cfg = {"workload": f"BASELINE configs[1]: {copies} x CylinderFlow mesh (1923 nodes, 11070 directed edges) per GPU, "
                       "block-diagonal batch; MGN processor (only_processor=True), H=128, 15 steps, fwd+bwd",
           "nodes_per_gpu": 1923 * copies, "edges_per_gpu": 11070 * copies, "hidden": H, "mp_steps": STEPS_MP,
           "parallelism": f"dp{n_gpus}",
           "numerics": "fp16 tensor-core operands, fp32 accumulation; residual adds, RMSNorm, aggregation and gradients in "
                       "fp32; edge stream held between steps as fp16 image + fp16 remainder (>= 21 bits)",
           "l2": "inputs larger than L2: every step streams a 181 MB edge-latent version (91 MB image + 91 MB remainder) "
                 "and a 181 MB edge-gradient version; 3.5 GB of saved state per fwd+bwd (L2 126 MB)"}
    if copies_note:
        cfg["sample"] = copies_note
    return cfg


# ------------------------------------------------------------------ B200 arm
def run_b200(args):
    import torch.distributed as dist

    from graph_physics_b200 import _lib as L
    from graph_physics_b200 import graph as G
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

    prog.phase("set-up: mesh batch, model, pinned inputs")
    copies = args.copies
    ei1, n1 = load_mesh()
    ei = G.tile_batch(ei1.to(dev), n1, copies)
    N, E = n1 * copies, ei.shape[1]
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
        proc.EVENTS = {"edge_bwd": []} if with_events else None
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

    prog.phase("csr build + eager pass (per-launch events of edge_bwd)")
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

    if rank == 0:
        hbm, tf, which = peaks()
        ms_step = ms / args.steps
        value = world * E / (ms_step * 1e-3)
        durs = [a.elapsed_time(b) for a, b in ev["edge_bwd"]]
        edge_bwd_ms = float(np.mean(durs)) if durs else float("nan")
        alg_bytes = E * EDGE_BWD_BYTES_EDGE + N * EDGE_BWD_BYTES_NODE
        ach = alg_bytes / (edge_bwd_ms * 1e-3) / 1e9
        step_bytes = STEPS_MP * (E * BYTES_EDGE_STEP + N * BYTES_NODE_STEP)
        step_flops = STEPS_MP * (E * FLOPS_EDGE_STEP + N * FLOPS_NODE_STEP)
        line = {
            "metric": METRIC, "value": value, "unit": "edges/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16",
            "data": "synthetic", "config": workload_config(world, copies),
            "e2e": {"value": world * E / (ms_e2e / e2e_steps * 1e-3), "unit": "edges/s",
                    "h2d_bytes_per_step": int(x0_h.numel() + e0_h.numel() + gx_h.numel()) * 4,
                    "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / e2e_steps},
            "gpu_launches": launches,
            "launch_mode": "cuda_graph_replay" if graphed else "eager",
            "ms_per_step_eager": ms_eager / args.steps,
            "forward_only": {"value": world * E / (ms_fwd / args.steps * 1e-3), "unit": "edges/s",
                             "ms_per_step": ms_fwd / args.steps, "what": "15 blocks forward under no_grad"},
            "clocks": clk,
            "roofline": {"bound": "hbm", "kernel": "edge_bwd_kernel", "achieved": ach, "peak": hbm, "unit": "GB/s",
                         "frac": ach / hbm, "traffic": EDGE_BWD_DRAM_BYTES_NCU if copies == 32 else None,
                         "traffic_source": "ncu --set full, profiles/r01_ncu_edge_bwd_v6.md (bytes per launch)",
                         "peak_source": which,
                         "ms_per_launch": edge_bwd_ms, "launches_timed": len(durs),
                         "timed_in": "eager pass of the same steps (one CUDA event pair per launch) run before the "
                                     "graph-replayed region of `value`",
                         "algorithmic_bytes_per_launch": alg_bytes},
            "step_roofline": {"hbm_frac": step_bytes / (ms_step * 1e-3) / 1e9 / hbm,
                              "tensor_frac": step_flops / (ms_step * 1e-3) / 1e12 / tf,
                              "algorithmic_GB_per_step": step_bytes / 1e9, "algorithmic_TFLOP_per_step": step_flops / 1e12,
                              "edge_bwd_share_of_step": edge_bwd_ms * STEPS_MP / ms_step},
        }
        # the GPU numbers exist from here on: a stall in the baseline legs below cannot take them along (the watchdog
        # prints this line); stderr carries it at once
        prog.partial = line
        print("[bench] result before the baseline legs: " + json.dumps(line), file=sys.stderr, flush=True)
        if args.cpu_seconds > 0 and world == 1:      # the CPU leg runs at N=1 only (rank 0 would hold N-1 ranks at a barrier)
            prog.phase("cpu_baseline leg (oracle port on the host cores)")
            threads = os.cpu_count() or 1
            eps, t, e_cpu = cpu_oracle_run(args.cpu_copies, 2, 1, threads)
            line["cpu_baseline"] = {"value": eps, "unit": "edges/s", "cores": threads, "kind": "port",
                                    "sample": f"{args.cpu_copies} of {copies} meshes ({e_cpu} edges), 15 steps fwd+bwd, "
                                              f"fp32 oracle port, median of 2 after 1 warm-up, {t:.2f} s per pass"}
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
        if os.environ.get("MGN_BENCH_BREAKDOWN"):       # developer aid: per-kernel device time of one extra step (stderr)
            names = ["pack", "p_fwd", "edge_fwd", "node_fwd", "node_bwd", "node_wgrad1", "edge_bwd", "edge_wgrad1", "dp_gather", "p_bwd"]
            proc.EVENTS = {n: [] for n in names}
            s0, e0_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            step_resident()
            e0_.record()
            torch.cuda.synchronize()
            evs, proc.EVENTS = proc.EVENTS, None
            tot = 0.0
            for n in names:
                d = [a.elapsed_time(b) for a, b in evs[n]]
                tot += sum(d)
                print(f"  {n:12s} n={len(d):3d} mean {1e3 * sum(d) / max(len(d), 1):8.1f} us  total {sum(d):7.3f} ms",
                      file=sys.stderr)
            print(f"  kernels {tot:.3f} ms of step {s0.elapsed_time(e0_):.3f} ms", file=sys.stderr)
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
    ap.add_argument("--copies", type=int, default=32, help="meshes per GPU")
    ap.add_argument("--cpu-copies", type=int, default=1, help="meshes in the CPU sample")
    ap.add_argument("--cpu-seconds", type=float, default=1.0, help="0 disables the cpu_baseline leg")
    ap.add_argument("--no-cuda-graph", action="store_true", help="time eager launches instead of a graph replay")
    ap.add_argument("--phase-timeout", type=float, default=240.0,
                    help="seconds one phase may take before the watchdog reports it and exits 3 (0 = off)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
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
