"""Drop-in `graphphysics.models.processors.EncodeProcessDecode` (reference
graphphysics/models/processors.py:27-137): same constructor arguments, attributes
(`K`, `d`, `temperature`, `hidden_size`, `only_processor`, `use_diagonal`,
`nodes_encoder`, `edges_encoder`, `decode_module`, `processor_list`) and state-dict keys,
so `get_model` (training/parse_parameters.py:91-101), `Simulator` (models/simulator.py:179,
327, 332-342) and the Lightning module pick it up unchanged.

`forward(graph)` reads `graph.x`, `graph.edge_attr`, `graph.edge_index` only (duck typed: a
torch_geometric `Data` / `Batch` or any object with those attributes).  The 15-block hot loop
(`processors.py:130-131`) runs as one fused autograd node on the sm_100a kernels
(graph_physics_b200.processor); the two encoder MLPs run on the fused node kernels
(graph_physics_b200.encoder) when their shape allows (input width <= 128) and
`fused_encoders` is left on; the decoder (no RMSNorm, 2-3 outputs) runs on the same kernels with their no-norm flag
(last layer zero-padded to 128 rows).
"""
from __future__ import annotations

from types import SimpleNamespace

import torch
import torch.nn as nn

from .. import encoder as _enc
from .. import processor as _proc
from .layers import GraphNetBlock, block_param_list, build_mlp


class Graph(SimpleNamespace):
    """Minimal stand-in for torch_geometric.data.Data where PyG is not installed."""

    def to(self, device):
        return Graph(**{k: (v.to(device) if torch.is_tensor(v) else v) for k, v in self.__dict__.items()})

    def clone(self):
        return Graph(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in self.__dict__.items()})


class EncodeProcessDecode(nn.Module):
    fused_encoders = True      # set False (class or instance) to run the encoders as eager torch modules

    def __init__(self, message_passing_num: int, node_input_size: int, edge_input_size: int, output_size: int,
                 hidden_size: int = 128, only_processor: bool = False, num_mixture_components: int = 0,
                 temperature: float = None, use_diagonal: bool = True):
        super().__init__()
        self.only_processor = only_processor
        self.hidden_size = hidden_size
        self.use_diagonal = use_diagonal
        self.d = output_size
        self.K = num_mixture_components
        self.temperature = temperature
        if not self.only_processor:
            self.nodes_encoder = build_mlp(in_size=node_input_size, hidden_size=hidden_size, out_size=hidden_size)
            self.edges_encoder = build_mlp(in_size=edge_input_size, hidden_size=hidden_size, out_size=hidden_size)
            if num_mixture_components == 0:
                self.decode_module = build_mlp(in_size=hidden_size, hidden_size=hidden_size, out_size=output_size,
                                               layer_norm=False)
            else:
                # GMM decoder heads (reference layers.py:116-195) are decoder-side and outside the
                # processor path this package accelerates ("OUT OF SCOPE").
                raise NotImplementedError("graph_physics_b200: mixture-density decoder heads "
                                          "(num_mixture_components > 0) are not part of the MGN processor path")
        self.processor_list = nn.ModuleList(
            [GraphNetBlock(hidden_size=hidden_size) for _ in range(message_passing_num)])

    def _encode(self, mlp, feat):
        if self.fused_encoders and feat.is_cuda and feat.dim() == 2 and _enc.supported(mlp, feat.shape[1]):
            return _enc.run_encoder(mlp, feat)
        return mlp(feat)

    def forward(self, graph) -> torch.Tensor:
        edge_index = graph.edge_index
        if self.only_processor:
            x, edge_attr = graph.x, graph.edge_attr
        else:
            x = self._encode(self.nodes_encoder, graph.x)
            edge_attr = self._encode(self.edges_encoder, graph.edge_attr)
        if len(self.processor_list) > 0:
            for block in self.processor_list:
                block.check_supported()
            # the reference discards the final edge latents (processors.py:130-137): no exit permutation
            x, _ = _proc.run_processor(x, edge_index, edge_attr, [block_param_list(b) for b in self.processor_list],
                                       need_edges=False)
        if self.only_processor:
            return x
        if self.fused_encoders and x.is_cuda and _enc.decoder_supported(self.decode_module):
            return _enc.run_decoder(self.decode_module, x)
        return self.decode_module(x)
