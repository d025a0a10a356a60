"""ORACLE / TEST INFRASTRUCTURE ONLY -- seeded synthetic processor inputs
(SURVEY.md section 8(d) "Synthetic inputs": x0, e0 ~ RMSNorm(N(0,1)), the
distribution both encoders produce since they end in RMSNorm with scale 1;
upstream gradient ~ N(0,1)/sqrt(N*H)).  Used by tests/golden/make_golden.py and by
the tests that compare against the committed fixtures, so both sides see the
same bits.  torch's CPU generator is deterministic for a fixed torch build; the
GPU box runs the same image."""
from __future__ import annotations

import torch

from .mgn_oracle import rms_norm


def processor_inputs(num_nodes: int, num_edges: int, hidden: int = 128, seed: int = 0,
                     dtype=torch.float32):
    g = torch.Generator().manual_seed(1000 + seed)
    ones = torch.ones(hidden, dtype=torch.float64)
    x0 = rms_norm(torch.randn(num_nodes, hidden, generator=g, dtype=torch.float64), ones)
    e0 = rms_norm(torch.randn(num_edges, hidden, generator=g, dtype=torch.float64), ones)
    gx = torch.randn(num_nodes, hidden, generator=g, dtype=torch.float64) / (num_nodes * hidden) ** 0.5
    return x0.to(dtype), e0.to(dtype), gx.to(dtype)
