"""graph_physics_b200 -- B200 (sm_100a) implementation of graph-physics' MeshGraphNet
message-passing processor behind the reference's `graphphysics.models` API.

    from graph_physics_b200.models import EncodeProcessDecode, GraphNetBlock, build_mlp, RMSNorm
    from graph_physics_b200 import graph      # device-side graph set-up (edges, CSR, K-hop, MeshMask)

The compute path is libmgn_b200.so (C ABI: include/mgn_b200.h, built in-tree by
`python -m graph_physics_b200.build`).  There is no CPU / eager fallback.
"""
__version__ = "0.1.0"
