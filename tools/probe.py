#!/usr/bin/env python
"""Runs csrc/probe.cu on the current GPU and prints the SM-local constants the kernel design leans on."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_physics_b200 import _lib as L  # noqa: E402

L.require_device()
iters = 256
out = torch.zeros(64, dtype=torch.int64, device="cuda")
for _ in range(2):
    L.check(L.lib().mgn_probe_sm(L.p(out), L.i32(iters), L.stream()))
    torch.cuda.synchronize()
o = out.tolist()
np_ = iters // 8 + 1
print(f"tcgen05.ld.32x32b.x32 (+wait), cycles per load per warp: nw=1 {o[0]/iters:.1f}  nw=4 {o[1]/iters:.1f}  "
      f"nw=8 {o[2]/iters:.1f}  nw=16 {o[3]/iters:.1f}")
print(f"  => 128x128 fp32 accumulator read by 16 warps (one x32 each): {o[3]/iters:.0f} cycles; "
      f"TMEM read B/clk/SM at nw=16: {16*4096*iters/o[3]:.0f}")
print(f"4 loads in flight, cycles per load per warp: nw=4 {o[4]/iters:.1f}  nw=16 {o[5]/iters:.1f} "
      f"(B/clk/SM nw=16: {16*4096*iters/o[5]:.0f})")
print(f"tcgen05.st.32x32b.x32, nw=16: {o[6]/iters:.1f} cycles per store per warp (B/clk/SM {16*4096*iters/o[6]:.0f})")
print(f"product M128 N128 K128 issue->barrier: 1 product {o[7]/np_:.0f}  2 back-to-back {o[8]/np_:.0f}  4 {o[9]/np_:.0f} cycles")
print(f"product M64 N128 K128: {o[10]/np_:.0f} cycles;  M128 N256 K128: {o[11]/np_:.0f} cycles")
print(f"product with 12 warps reading TMEM: {o[12]/np_:.0f} cycles per product; readers' phase {o[13]} cycles "
      f"({o[13]/iters:.1f} per load)")
print(f"one product, issuing thread's cycles in 8 x tcgen05.mma / tcgen05.commit / mbarrier wait: alone {o[14]/np_:.0f} / {o[15]/np_:.0f} / "
      f"{o[16]/np_:.0f};  with 12 warps streaming shared memory {o[17]/np_:.0f} / {o[18]/np_:.0f} / {o[19]/np_:.0f}")

# ---- SM <-> L2 / DRAM path with bulk copies (mgn_probe_mem)
if "--no-mem" not in sys.argv:
    grid, chunks = 148, 256
    buf = torch.zeros(grid * chunks * 32768 + 131072, dtype=torch.uint8, device="cuda")
    table = torch.zeros(65536, dtype=torch.uint8, device="cuda")
    mo = torch.zeros(grid * 4, dtype=torch.int64, device="cuda")
    names = {0: "g2s, L2-resident source (all CTAs read the same 128 KB)", 1: "g2s, streaming from DRAM",
             2: "s2g, streaming", 3: "reduce-add f32, streaming", 4: "g2s + s2g streaming together",
             5: "s2g streaming + dependent 16-byte L2 loads from another warp", 6: "dependent 16-byte L2 loads alone",
             7: "g2s + s2g streaming (2 + 2 slots) while another thread runs 32 products", 8: "the 32 products alone", 9: "32 x (64 KB reduce-add, then a product, same thread)",
             10: "32 x (64 KB bulk store, then a product, same thread)"}
    for mode in range(11):
        best = None
        for _ in range(3):
            mo.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            L.check(L.lib().mgn_probe_mem(L.p(buf), L.p(table), L.p(mo), L.i32(mode), L.i32(chunks), L.i32(grid), L.stream()))
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b)
            if best is None or ms < best[0]:
                best = (ms, mo.view(grid, 4).clone())
        ms, r = best
        cyc = r[:, 0].double()
        moved = grid * chunks * 32768 * (2 if mode in (4, 7) else 1)
        line = f"mode {mode} ({names[mode]}): {ms * 1e3:.0f} us"
        if mode >= 8:
            pass
        elif mode != 6:
            line += (f", {moved / ms / 1e6:.0f} GB/s aggregate, {chunks * 32768 * (2 if mode in (4, 7) else 1) / cyc.mean().item():.1f} B/clk/SM "
                     f"(CTA cycles mean {cyc.mean().item():.0f} max {cyc.max().item():.0f})")
        if mode >= 7:
            line += (f"; per product: issue + commit {r[:, 2].double().mean().item() / 32:.0f} cycles, mbarrier wait "
                     f"{r[:, 3].double().mean().item() / 32:.0f} cycles")
        elif mode >= 5:
            lat = r[:, 2].double().sum().item() / max(r[:, 3].double().sum().item(), 1)
            line += f"; load latency {lat:.0f} cycles (n={int(r[:, 3].sum().item())})"
        print(line)
