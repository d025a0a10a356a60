"""Drop-in `graphphysics.models.simulator.Simulator` (reference graphphysics/models/simulator.py:128-405) for the
deterministic MeshGraphNet path (`model.K == 0`): the steps either side of the network in a training / validation /
rollout step.

    target_delta_normalized = output_normalizer(y - x[:, out_start:out_end])           simulator.py:215-235
    node features           = node_normalizer([x[:, f_start:f_end] | one_hot(type, 9)]) simulator.py:237-268, 284-288
    edge features           = edge_normalizer(edge_attr)                               simulator.py:290-293
    network_output          = model(graph)                                             simulator.py:327
    eval only:  outputs     = x[:, out_start:out_end] + output_normalizer.inverse(network_output)   :304-307, 332-336

Same constructor arguments, attribute names, normaliser sub-modules (state-dict keys of a reference checkpoint) and
return triple.  What differs is how it runs on the device: the one-hot block is written by a comparison against a
class range straight into the feature matrix (no int64 one-hot tensor, no concatenation of mixed dtypes), the
normalisers accumulate without reading a counter back to the host (`layers.Normalizer`), so the whole pre-processing is
a fixed sequence of device operations and a training step `Simulator -> loss` can be captured in one CUDA graph
together with the fused kernels.  The graph handed to the model is `processors.Graph` (the duck type the models of this
package read: `.x`, `.edge_attr`, `.edge_index`, `.pos`).

The mixture-density sampling branch (`model.K > 0`, simulator.py:337-349) belongs to the GMM decoder heads, outside
this package's path: it raises.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
import torch.nn as nn

from ..utils.nodetype import NodeType
from .layers import Normalizer
from .processors import Graph


class Simulator(nn.Module):
    check_node_types = True     # device-side range assert of the node-type column (see _get_one_hot_type)

    def __init__(self, node_input_size: int, edge_input_size: int, output_size: int, feature_index_start: int,
                 feature_index_end: int, output_index_start: int, output_index_end: int, node_type_index: int,
                 model: nn.Module, device, model_dir: str = "checkpoint/simulator.pth"):
        super().__init__()
        self.node_input_size = node_input_size
        self.edge_input_size = edge_input_size if edge_input_size > 0 else None
        self.output_size = output_size
        self.feature_index_start = feature_index_start
        self.feature_index_end = feature_index_end
        self.node_type_index = node_type_index
        self.output_index_start = output_index_start
        self.output_index_end = output_index_end
        self.model_dir = model_dir
        self.model = model.to(device)
        self._output_normalizer = Normalizer(size=output_size, name="output_normalizer", device=device)
        self._node_normalizer = Normalizer(size=node_input_size, name="node_normalizer", device=device)
        self._edge_normalizer = (Normalizer(size=edge_input_size, name="edge_normalizer", device=device)
                                 if self.edge_input_size is not None else None)
        self.device = device

    # ---- pieces, with the reference's names -------------------------------------------------------------------
    def _get_pre_target(self, inputs) -> torch.Tensor:
        return inputs.x[:, self.output_index_start:self.output_index_end]

    def _get_target_normalized(self, inputs, is_training: bool = True) -> torch.Tensor:
        return self._output_normalizer(inputs.y - self._get_pre_target(inputs), is_training)

    def _get_one_hot_type(self, inputs) -> torch.Tensor:
        """[N, NodeType.SIZE] one-hot of the node-type column, in the features' dtype (the reference makes an int64
        one-hot that the concatenation promotes: same values)."""
        node_type = inputs.x[:, self.node_type_index].long()
        if self.check_node_types and node_type.is_cuda:
            # the reference's torch.nn.functional.one_hot raises for a class outside [0, SIZE); here the check runs on
            # the device (no host read: the step stays capturable) and a violation surfaces as a CUDA device assert
            torch._assert_async(((node_type >= 0) & (node_type < int(NodeType.SIZE))).all())
        classes = torch.arange(int(NodeType.SIZE), device=node_type.device)
        return (node_type.unsqueeze(-1) == classes).to(inputs.x.dtype)

    def _build_node_features(self, inputs, one_hot_type: torch.Tensor) -> torch.Tensor:
        features = inputs.x[:, self.feature_index_start:self.feature_index_end]
        return torch.cat([features, one_hot_type.to(features.dtype)], dim=1)

    def _build_input_graph(self, inputs, is_training: bool):
        target_delta_normalized = self._get_target_normalized(inputs, is_training)
        node_features = self._build_node_features(inputs, self._get_one_hot_type(inputs))
        node_features_normalized = self._node_normalizer(node_features, is_training)
        if self._edge_normalizer is not None:
            edge_attr = self._edge_normalizer(inputs.edge_attr, is_training)
        else:
            edge_attr = inputs.edge_attr
        graph = Graph(x=node_features_normalized, pos=getattr(inputs, "pos", None), edge_attr=edge_attr,
                      edge_index=inputs.edge_index)
        return graph, target_delta_normalized

    def _build_outputs(self, inputs, network_output: torch.Tensor) -> torch.Tensor:
        return self._get_pre_target(inputs) + self._output_normalizer.inverse(network_output)

    # ---- the call the Lightning module makes (lightning_module.py:109-123, 183) ------------------------------------
    def forward(self, inputs) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
        graph, target_delta_normalized = self._build_input_graph(inputs=inputs, is_training=self.training)
        network_output = self.model(graph)
        if self.training:
            return network_output, target_delta_normalized, None
        if getattr(self.model, "K", 0) != 0:
            raise NotImplementedError("graph_physics_b200: mixture-density sampling (model.K > 0) is not part of the "
                                      "MeshGraphNet processor path")
        return network_output, target_delta_normalized, self._build_outputs(inputs, network_output)

    @torch.no_grad()
    def predict(self, inputs) -> torch.Tensor:
        """The `outputs` of an evaluation-mode `forward` for a graph that has no target (`inputs.y` is not read and no
        normaliser accumulates): what a rollout step needs -- `rollout.Rollout(predict_fn=simulator.predict, ...)`
        replays it from one CUDA graph per time step."""
        if getattr(self.model, "K", 0) != 0:
            raise NotImplementedError("graph_physics_b200: mixture-density sampling (model.K > 0) is not part of the "
                                      "MeshGraphNet processor path")
        node_features = self._build_node_features(inputs, self._get_one_hot_type(inputs))
        edge_attr = inputs.edge_attr
        if self._edge_normalizer is not None:
            edge_attr = self._edge_normalizer(edge_attr, False)
        graph = Graph(x=self._node_normalizer(node_features, False), pos=getattr(inputs, "pos", None),
                      edge_attr=edge_attr, edge_index=inputs.edge_index)
        return self._build_outputs(inputs, self.model(graph))

    def freeze_all(self) -> None:
        for param in self.model.parameters():
            param.requires_grad = False

    # ---- checkpoints: the reference's file layout (simulator.py:359-405) ------------------------------------------
    _NORMALIZER_KEYS = ("_output_normalizer", "_node_normalizer", "_edge_normalizer")

    def load_checkpoint(self, ckpdir: Optional[str] = None) -> None:
        ckpdir = self.model_dir if ckpdir is None else ckpdir
        checkpoint = torch.load(ckpdir, map_location=self.device, weights_only=False)
        self.load_state_dict(checkpoint["model"])
        for key in self._NORMALIZER_KEYS:
            state = checkpoint.get(key) or {}
            normalizer = getattr(self, key, None)
            if normalizer is not None and state:
                for attr_name, value in state.items():
                    if torch.is_tensor(value) and attr_name in normalizer._buffers:
                        normalizer._buffers[attr_name].copy_(value)        # keeps addresses a captured graph holds
                    else:
                        setattr(normalizer, attr_name, value)

    def save_checkpoint(self, savedir: Optional[str] = None) -> None:
        savedir = self.model_dir if savedir is None else savedir
        if os.path.dirname(savedir):
            os.makedirs(os.path.dirname(savedir), exist_ok=True)
        to_save = {"model": self.state_dict()}
        for key in self._NORMALIZER_KEYS:
            normalizer = getattr(self, key, None)
            to_save[key] = normalizer.get_variable() if normalizer is not None else None
        torch.save(to_save, savedir)
