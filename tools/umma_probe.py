"""Diagnostic (GPU): for every self-test case print the error with the default
descriptor strides and, when that fails, search LBO/SBO alternatives."""
import itertools
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gpu_umma import CASES, run_case

cands = [-1, 16, 128, 1024, 2048, 16384]
for name, cfg in CASES.items():
    err = run_case(cfg)
    print(f"{name}: default err={err:.3e}", flush=True)
    if err < 2e-3:
        continue
    for la, sa, lb, sb in itertools.product(cands, cands, cands, cands):
        if not cfg[2] and (la != -1 or sa != -1):
            continue
        if (not cfg[3] or cfg[7]) and (lb != -1 or sb != -1):
            continue
        e = run_case(cfg, la, sa, lb, sb)
        if e < 2e-3:
            print(f"   OK with lbo_a={la} sbo_a={sa} lbo_b={lb} sbo_b={sb} err={e:.3e}", flush=True)
