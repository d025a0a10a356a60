"""Mirror of the `graphphysics.utils` names the MeshGraphNet training / rollout path uses."""
from .loss import L1SmoothLoss, L2Loss  # noqa: F401
from .nodetype import NodeType  # noqa: F401
