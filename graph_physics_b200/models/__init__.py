"""Mirror of the `graphphysics.models` surface for the MeshGraphNet path."""
from .layers import GraphNetBlock, Normalizer, RMSNorm, build_mlp  # noqa: F401
from .processors import EncodeProcessDecode, Graph  # noqa: F401
from .simulator import Simulator  # noqa: F401
