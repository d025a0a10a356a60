"""The two step functions of the reference's Lightning module as plain functions (no trainer, optimiser or logging: the
control plane stays the caller's): `training_step` (training/lightning_module.py:112-123) and the rollout of a
validation / prediction trajectory (`_make_prediction`, :168-202) on `rollout.Rollout` with a `Simulator`.

Both are fixed sequences of device operations (the Simulator, normalisers and losses of this package read nothing back
on the host), so `processor.CapturedStep(lambda: training_step(...).backward())` captures a whole training step and
`Rollout` replays one CUDA graph per time step.
"""
from __future__ import annotations

from typing import Iterable, Optional, Sequence, Tuple

import torch


def training_step(simulator, batch, loss_fn, loss_masks: Sequence[int]) -> torch.Tensor:
    """loss = loss_fn(target_delta_normalized, network_output, node_type, masks) for a `Simulator` in training mode."""
    node_type = batch.x[:, simulator.node_type_index]
    network_output, target_delta_normalized, _ = simulator(batch)
    return loss_fn(target_delta_normalized, network_output, node_type, masks=loss_masks)


def make_rollout(simulator, graph, prev_slice: Optional[Tuple[int, int]] = None, capture: bool = True):
    """`rollout.Rollout` wired to a Simulator: predictions = `simulator.predict` (evaluation outputs), the fed-back
    columns are the simulator's output range, boundary nodes come from its node-type column."""
    from .rollout import Rollout
    simulator.eval()
    return Rollout(simulator.predict, graph, out_slice=(simulator.output_index_start, simulator.output_index_end),
                   node_type_index=simulator.node_type_index, prev_slice=prev_slice, capture=capture)


def rollout_rmse(predictions: torch.Tensor, targets: Iterable[torch.Tensor]) -> torch.Tensor:
    """Per-step RMSE of a rollout over all nodes and components ([T]; the quantity `validation_step` accumulates,
    lightning_module.py:204-232), computed on the device."""
    t = torch.stack(list(targets)) if not torch.is_tensor(targets) else targets
    return torch.sqrt(torch.mean((predictions - t.to(predictions.device)) ** 2, dim=(1, 2)))
