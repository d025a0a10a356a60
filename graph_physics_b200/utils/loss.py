"""Masked node losses of the reference's training step (graphphysics/utils/loss.py:10-112; `training_step`,
training/lightning_module.py:112-123): the mean of the per-component error over the nodes whose type is in `masks`
and, optionally, which are not in `selected_indexes`.

The reference selects the rows with a boolean index (`errors[mask]`), whose result size is data dependent: a device ->
host read in the middle of the step, and a step that cannot be captured in a CUDA graph.  Here the mask multiplies the
errors and the mean is `sum / (count * d)` -- the same number up to the order of an fp32 sum, from a fixed sequence of
device operations.  An empty selection gives NaN, as `torch.mean` of an empty tensor does in the reference.

The mixture-density losses (loss.py:115-327) belong to the GMM decoder heads, which are outside this package's path.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn.functional as F
from torch.nn.modules.loss import _Loss


def _prepare_mask_for_loss(network_output: torch.Tensor, node_type: torch.Tensor, masks: Sequence[int],
                           selected_indexes: Optional[torch.Tensor] = None) -> torch.Tensor:
    """loss.py:10-26, on the device of `node_type` (the reference builds the index range on the host)."""
    mask = node_type == int(masks[0])
    for m in masks[1:]:
        mask = torch.logical_or(mask, node_type == int(m))
    if selected_indexes is not None:
        keep = torch.ones(network_output.shape[0], dtype=torch.bool, device=node_type.device)
        keep[selected_indexes.to(node_type.device).long()] = False
        mask = torch.logical_and(keep, mask)
    return mask


def _masked_mean(errors: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    m = mask.unsqueeze(-1)
    kept = torch.where(m, errors, torch.zeros((), dtype=errors.dtype, device=errors.device))   # (no 0 * inf)
    return kept.sum() / (mask.sum().to(errors.dtype) * errors.shape[-1])


class L2Loss(_Loss):
    @property
    def __name__(self):
        return "MSE"

    def forward(self, target, network_output, node_type, masks, selected_indexes=None):
        mask = _prepare_mask_for_loss(network_output, node_type, masks, selected_indexes)
        return _masked_mean((network_output - target) ** 2, mask)


class L1SmoothLoss(_Loss):
    def __init__(self, beta: float = 1.0, **kwargs):
        super().__init__(**kwargs)
        self.beta = beta

    @property
    def __name__(self):
        return "L1Smooth"

    def forward(self, target, network_output, node_type, masks, selected_indexes=None):
        mask = _prepare_mask_for_loss(network_output, node_type, masks, selected_indexes)
        return _masked_mean(F.smooth_l1_loss(network_output, target, reduction="none", beta=self.beta), mask)
