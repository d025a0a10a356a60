"""GPU: the CSR of a collated batch assembled from the samples' device-built CSRs (`batch.concat_csr`) is bit-identical
to `mgn_build_csr` run on the batch's own `edge_index`, and the processor picks the registered CSR up.

(Written after the round's GPU budget was spent: the logic is device independent and covered by tests/test_cpu_batch.py
against the oracle; this file sorts last so that it cannot hide the results of the kernel tests.)"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_concat_csr_equals_device_build(golden_dir):
    from graph_physics_b200 import batch as B, graph as G, processor as proc
    from graph_physics_b200.models import Graph
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    dev = "cuda"
    ei_mesh = torch.as_tensor(z["edge_index"].astype(np.int64), device=dev)
    n_mesh = int(z["points"].shape[0])
    g = torch.Generator(device="cpu").manual_seed(9)
    samples = []
    for n, ei in ((n_mesh, ei_mesh), (300, torch.randint(0, 300, (2, 1000), generator=g).to(dev)), (n_mesh, ei_mesh)):
        samples.append(Graph(x=torch.randn(n, 128, generator=g).to(dev),
                             edge_attr=torch.randn(ei.shape[1], 128, generator=g).to(dev), edge_index=ei))
    csrs = [proc.csr_for(s.edge_index, s.x.shape[0]) for s in samples]
    b = B.collate(samples, csrs=csrs)
    want = G.build_csr(b.edge_index, b.num_nodes)
    for name in ("rowptr", "perm", "senders", "receivers"):
        assert torch.equal(getattr(b.csr, name), getattr(want, name)), name
    assert b.csr.identity == want.identity
    assert proc.csr_for(b.edge_index, b.num_nodes) is b.csr
