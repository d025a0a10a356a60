"""GPU: the partitioned processor (graph_physics_b200.dist_processor) against the un-partitioned one on the same
device -- several partitions of one mesh stepped in lock-step with copies in place of the collectives -- and against
the fp64 oracle.  The two GPU paths run the same kernels on the same rows; they differ only in the composition of the
128-row tiles (hence in summation order), so they agree far inside the path's tolerance."""
import os

import numpy as np
import pytest
import torch

from oracle import mgn_oracle as mo
from oracle.synth import processor_inputs


def block_param_list(blk):
    from graph_physics_b200.models.layers import block_param_list as f
    return f(blk)

pytestmark = pytest.mark.gpu


def _setup(ei, n, steps, seed):
    from graph_physics_b200.models import EncodeProcessDecode
    p = mo.init_params(steps, hidden=128, seed=seed)
    model = EncodeProcessDecode(steps, 128, 128, 128, hidden_size=128, only_processor=True).cuda()
    model.load_state_dict({k: v.clone() for k, v in p.items()})
    x0, e0, gx = processor_inputs(n, ei.shape[1], 128, seed=seed + 1)
    return model, p, x0, e0, gx


def _whole(model, x0, ei, e0, gx, ge):
    from graph_physics_b200 import processor as proc
    xd, ed = x0.cuda().requires_grad_(True), e0.cuda().requires_grad_(True)
    for p in model.parameters():
        p.grad = None
    blocks = [block_param_list(blk) for blk in model.processor_list]
    xo, eo = proc.run_processor(xd, ei.cuda(), ed, blocks)
    torch.autograd.backward([xo, eo], [gx.cuda(), ge.cuda()])
    return xo.detach(), eo.detach(), xd.grad, ed.grad, proc.LAST_GRAD_BLOCKS.clone()


@pytest.mark.parametrize("world", [2, 5])
def test_partitions_match_whole_mesh(golden_dir, world):
    from graph_physics_b200 import dist_processor as DP
    from graph_physics_b200 import partition as PT
    z = np.load(os.path.join(golden_dir, "cylinder_mesh.npz"))
    n = int(z["points"].shape[0])
    ei = torch.as_tensor(z["edge_index"].astype(np.int64))
    perm = torch.randperm(ei.shape[1], generator=torch.Generator().manual_seed(1))
    ei = ei[:, perm]                                   # not receiver-sorted: exercises the local permutation
    steps = 3
    model, p, x0, e0, gx = _setup(ei, n, steps, 20 + world)
    ge = 0.5 * processor_inputs(n, ei.shape[1], 128, seed=99)[1] / (ei.shape[1] ** 0.5)
    xo, eo, gxd, ged, gw = _whole(model, x0, ei, e0, gx, ge)
    parts = [PT.build_partition(ei.cuda(), n, r, world) for r in range(world)]
    blocks = [block_param_list(blk) for blk in model.processor_list]
    res = DP.run_partitions_local(parts, x0.cuda(), e0.cuda(), blocks, gx.cuda(), ge.cuda())
    errs = {"x": mo.rel_l2(res["x_out"], xo), "e": mo.rel_l2(res["e_out"], eo),
            "gx": mo.rel_l2(res["g_x"], gxd), "ge": mo.rel_l2(res["g_e"], ged),
            "gw": mo.rel_l2(res["grads"], gw)}
    print(f"world={world}: halo sizes {[q.num_halo for q in parts]}, errs {errs}")
    assert errs["x"] < 3e-4 and errs["e"] < 3e-4, errs          # fp16 operand flips from the summation order only
    assert errs["gx"] < 2e-2 and errs["ge"] < 2e-2 and errs["gw"] < 2e-2, errs
    # and against the exact oracle, at the path's tolerance
    p64 = {k: v.double() for k, v in p.items()}
    ro, _ = mo.processor(x0.double(), ei, e0.double(), p64, steps)
    assert mo.rel_l2(res["x_out"].cpu(), ro) < 1e-3 * steps
    # inference mode (ping-pong buffers) gives the same forward
    res_inf = DP.run_partitions_local(parts, x0.cuda(), e0.cuda(), blocks)
    assert mo.rel_l2(res_inf["x_out"], res["x_out"]) < 1e-6 and mo.rel_l2(res_inf["e_out"], res["e_out"]) < 1e-6


def test_partitioned_tet_mesh(golden_dir):
    """The 3D family config 5 is made of (aneurysm fixture): 4 partitions, 2 steps, against the whole mesh."""
    from graph_physics_b200 import dist_processor as DP
    from graph_physics_b200 import graph as G
    from graph_physics_b200 import partition as PT
    z = np.load(os.path.join(golden_dir, "aneurysm_mesh.npz"))
    cells = torch.as_tensor(z["cells"].astype(np.int64)).cuda()
    n = int(z["points"].shape[0])
    ei = G.face_to_edge(G.cells_to_faces(cells), n).cpu()
    steps = 2
    model, p, x0, e0, gx = _setup(ei, n, steps, 31)
    ge = torch.zeros_like(e0)
    xo, eo, gxd, ged, gw = _whole(model, x0, ei, e0, gx, ge)
    parts = [PT.build_partition(ei.cuda(), n, r, 4) for r in range(4)]
    blocks = [block_param_list(blk) for blk in model.processor_list]
    res = DP.run_partitions_local(parts, x0.cuda(), e0.cuda(), blocks, gx.cuda(), None)
    errs = {"x": mo.rel_l2(res["x_out"], xo), "e": mo.rel_l2(res["e_out"], eo),
            "gx": mo.rel_l2(res["g_x"], gxd), "ge": mo.rel_l2(res["g_e"], ged),
            "gw": mo.rel_l2(res["grads"], gw)}
    print(f"tet, 4 partitions: edges {[int(q.edge_ids.numel()) for q in parts]}, halo {[q.num_halo for q in parts]}, "
          f"errs {errs}")
    assert errs["x"] < 3e-4 and errs["e"] < 3e-4, errs
    assert errs["gx"] < 2e-2 and errs["ge"] < 2e-2 and errs["gw"] < 2e-2, errs
