"""Drop-in `graphphysics.models.layers` surface for the MeshGraphNet path: RMSNorm,
build_mlp and GraphNetBlock with the reference's signatures, attribute names and
state-dict keys (graphphysics/models/layers.py:18-113, 630-746), without torch_geometric.

GraphNetBlock.forward runs on the fused sm_100a kernels (graph_physics_b200.processor);
it needs CUDA tensors on a B200 and raises otherwise -- there is no eager fallback.
RMSNorm / build_mlp are ordinary torch modules: they hold the parameters (so AdamW,
state_dict, Lightning checkpoints keep working) and serve the encoder / decoder MLPs,
which are outside the processor hot path.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import processor as _proc


class RMSNorm(nn.Module):
    """y = scale * x / (||x||_2 * d^-1/2 + eps)  (reference layers.py:18-74; eps is added to
    the RMS, no mean subtraction; `p` selects partial RMSNorm as in the reference)."""

    def __init__(self, d, p=-1.0, eps=1e-8, bias=False):
        super().__init__()
        self.eps = eps
        self.d = d
        self.p = p
        self.bias = bias
        self.scale = nn.Parameter(torch.ones(d))
        self.register_parameter("scale", self.scale)
        if self.bias:
            self.offset = nn.Parameter(torch.zeros(d))
            self.register_parameter("offset", self.offset)

    def forward(self, x):
        if self.p < 0.0 or self.p > 1.0:
            norm_x = x.norm(2, dim=-1, keepdim=True)
            d_x = self.d
        else:
            partial_size = int(self.d * self.p)
            partial_x, _ = torch.split(x, [partial_size, self.d - partial_size], dim=-1)
            norm_x = partial_x.norm(2, dim=-1, keepdim=True)
            d_x = partial_size
        rms_x = norm_x * d_x ** (-1.0 / 2)
        x_normed = x / (rms_x + self.eps)
        if self.bias:
            return self.scale * x_normed + self.offset
        return self.scale * x_normed


def build_mlp(in_size: int, hidden_size: int, out_size: int, nb_of_layers: int = 4,
              layer_norm: bool = True) -> nn.Module:
    """Linear, ReLU, (Linear, ReLU) * (L-2), Linear, [RMSNorm] as an nn.Sequential with the
    reference's module indices (layers.py:77-113)."""
    assert nb_of_layers >= 2, "The MLP must have at least 2 layers (input and output)."
    layers = [nn.Linear(in_size, hidden_size), nn.ReLU()]
    for _ in range(nb_of_layers - 2):
        layers.extend([nn.Linear(hidden_size, hidden_size), nn.ReLU()])
    layers.append(nn.Linear(hidden_size, out_size))
    module = nn.Sequential(*layers)
    if layer_norm:
        module = nn.Sequential(*module, RMSNorm(out_size))
    return module


class Normalizer(nn.Module):
    """Running-statistics feature normaliser of the reference (layers.py:265-393): same constructor arguments, buffers
    (`_acc_count`, `_num_accumulations`, `_acc_sum`, `_acc_sum_squared`: the state-dict keys of a reference
    checkpoint), `forward(batched_data, accumulate=True)`, `inverse` and `get_variable`.

    The reference decides `_num_accumulations < _max_accumulations` on the host, i.e. one device -> host read per call
    (three per training step through the Simulator), which also makes the step impossible to capture in a CUDA graph.
    Here the test stays on the device: the batch statistics are multiplied by the 0 / 1 outcome before they are added,
    so the accumulation is a fixed sequence of device operations with the same result."""

    def __init__(self, size: int, max_accumulations: int = 10 ** 5, std_epsilon: float = 1e-8,
                 name: str = "Normalizer", device="cuda"):
        super().__init__()
        self.name = name
        self.device = device
        self._max_accumulations = max_accumulations
        self._std_epsilon = torch.tensor(std_epsilon, dtype=torch.float32, requires_grad=False, device=device)
        self.register_buffer("_acc_count", torch.tensor(0.0, device=device))
        self.register_buffer("_num_accumulations", torch.tensor(0.0, device=device))
        self.register_buffer("_acc_sum", torch.zeros((1, size), dtype=torch.float32, device=device))
        self.register_buffer("_acc_sum_squared", torch.zeros((1, size), dtype=torch.float32, device=device))

    def forward(self, batched_data: torch.Tensor, accumulate: bool = True) -> torch.Tensor:
        if accumulate:
            self._accumulate(batched_data.detach())
        return (batched_data - self._mean()) / self._std_with_epsilon()

    def inverse(self, normalized_batch_data: torch.Tensor) -> torch.Tensor:
        return normalized_batch_data * self._std_with_epsilon() + self._mean()

    @torch.no_grad()
    def _accumulate(self, batched_data: torch.Tensor):
        # gate = 1 while fewer than max_accumulations batches were seen (layers.py:329-332), else 0: no host read
        gate = (self._num_accumulations < self._max_accumulations).to(torch.float32)
        self._acc_sum += gate * torch.sum(batched_data, dim=0, keepdim=True)
        self._acc_sum_squared += gate * torch.sum(batched_data ** 2, dim=0, keepdim=True)
        self._acc_count += gate * batched_data.shape[0]
        self._num_accumulations += gate

    def _mean(self) -> torch.Tensor:
        return self._acc_sum / torch.clamp(self._acc_count, min=1.0)

    def _std_with_epsilon(self) -> torch.Tensor:
        variance = self._acc_sum_squared / torch.clamp(self._acc_count, min=1.0) - self._mean() ** 2
        std = torch.sqrt(torch.clamp(variance, min=0.0))
        return torch.max(std, self._std_epsilon.to(std.device))

    def get_variable(self):
        return {"_max_accumulations": self._max_accumulations, "_std_epsilon": self._std_epsilon,
                "_acc_count": self._acc_count, "_num_accumulations": self._num_accumulations,
                "_acc_sum": self._acc_sum, "_acc_sum_squared": self._acc_sum_squared, "name": self.name}


KERNEL_H, KERNEL_LAYERS = 128, 4      # the shape the sm_100a kernels are built for


def _mlp_as_kernel_shape(mlp: nn.Sequential, in_blocks: int, hidden: int, nb_of_layers: int):
    """The 9 tensors (4 x weight, bias; scale) of a `build_mlp(in_blocks * hidden, hidden, hidden, nb_of_layers)` at
    the kernels' shape (hidden 128, four layers), built with differentiable torch ops so that autograd hands the
    gradients back to the module's own parameters:
      * hidden < 128: every matrix / vector is zero-padded; each `hidden`-wide column block of the first layer goes to
        the start of its own 128-wide block ([e | x_i | x_j] and [x | aggr] keep their block structure).  Padded
        latent columns stay exactly zero (zero rows, zero bias, zero scale), and RMSNorm divides by sqrt(hidden)
        (the kernels' `norm_dim`);
      * fewer than four layers: identity layers with zero bias are inserted before the last layer -- after a ReLU the
        activations are >= 0, so relu(I h + 0) = h exactly, forward and backward."""
    lin = [mlp[2 * i] for i in range(nb_of_layers)]
    norm = mlp[2 * nb_of_layers - 1]
    pad = KERNEL_H - hidden
    F = torch.nn.functional

    def pad_sq(w):          # [hidden, hidden] -> [128, 128]
        return F.pad(w, (0, pad, 0, pad)) if pad else w

    def pad_vec(v):
        return F.pad(v, (0, pad)) if pad else v
    w1 = lin[0].weight
    if pad:
        w1 = torch.cat([F.pad(w1[:, j * hidden:(j + 1) * hidden], (0, pad, 0, pad)) for j in range(in_blocks)], dim=1)
    out = [w1, pad_vec(lin[0].bias)]
    middles = [(pad_sq(l.weight), pad_vec(l.bias)) for l in lin[1:-1]]
    while len(middles) < KERNEL_LAYERS - 2:
        eye = torch.eye(KERNEL_H, dtype=w1.dtype, device=w1.device)
        middles.append((eye, torch.zeros(KERNEL_H, dtype=w1.dtype, device=w1.device)))
    for w, b in middles:
        out += [w, b]
    out += [pad_sq(lin[-1].weight), pad_vec(lin[-1].bias), pad_vec(norm.scale)]
    return out


def block_param_list(block: "GraphNetBlock"):
    """The 18 parameters of a block in ops.GRAD_LAYOUT (= state-dict) order, at the kernels' shape."""
    eb, nb = block.edge_block, block.node_block
    if block.hidden_size == KERNEL_H and block.nb_of_layers == KERNEL_LAYERS:
        return [eb[0].weight, eb[0].bias, eb[2].weight, eb[2].bias, eb[4].weight, eb[4].bias, eb[6].weight,
                eb[6].bias, eb[7].scale,
                nb[0].weight, nb[0].bias, nb[2].weight, nb[2].bias, nb[4].weight, nb[4].bias, nb[6].weight,
                nb[6].bias, nb[7].scale]
    return (_mlp_as_kernel_shape(eb, 3, block.hidden_size, block.nb_of_layers) +
            _mlp_as_kernel_shape(nb, 2, block.hidden_size, block.nb_of_layers))


class GraphNetBlock(nn.Module):
    """One message-passing step (reference layers.py:630-746):
       e_ = edge_block(cat[e, x[col], x[row]]); a = scatter-sum of e_ at col;
       x_ = node_block(cat[x, a]); returns (x + x_, e + e_), edges in the caller's order."""

    def __init__(self, hidden_size: int, nb_of_layers: int = 4, layer_norm: bool = True):
        super().__init__()
        self.hidden_size = hidden_size
        self.nb_of_layers = nb_of_layers
        self.layer_norm = layer_norm
        self.edge_block = build_mlp(3 * hidden_size, hidden_size, hidden_size, nb_of_layers, layer_norm)
        self.node_block = build_mlp(2 * hidden_size, hidden_size, hidden_size, nb_of_layers, layer_norm)

    def check_supported(self):
        """hidden_size <= 128 and nb_of_layers in 2..4 run on the 128-wide four-layer kernels (narrower / shallower
        blocks are zero-padded / get identity layers, `block_param_list`); layer_norm=False has no kernel."""
        if not (0 < self.hidden_size <= KERNEL_H) or not (2 <= self.nb_of_layers <= KERNEL_LAYERS) or not self.layer_norm:
            raise ValueError("graph_physics_b200.GraphNetBlock: the sm_100a kernels cover hidden_size <= 128, "
                             "nb_of_layers in {2, 3, 4}, layer_norm=True; there is no eager fallback")

    def forward(self, x, edge_index, edge_attr, size=None):
        self.check_supported()
        return _proc.run_processor(x, edge_index, edge_attr, [block_param_list(self)])
