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
