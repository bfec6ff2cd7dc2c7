"""Autograd + (optionally) ``DistributedDataParallel`` + ``torch.optim.AdamW``.

This is the step sequence of the reference (`src/llmtrain/training/trainer.py:362-395`):
zero grads → for each micro-batch: loss / accum, backward (inside ``no_sync()`` for all but the
last micro-batch under DDP) → ``clip_grad_norm_`` → ``optimizer.step()``.  Used for every
non-native combination (CPU/gloo in particular) and as the oracle the native engine is tested
against.
"""

from __future__ import annotations

import contextlib
from collections.abc import Callable
from typing import Any

import torch
from torch import nn
from torch.nn.parallel import DistributedDataParallel

from llmtrain.engine.base import Batch, TrainEngine


def move_batch(batch: Batch, device: torch.device) -> Batch:
    """Tensor values to ``device`` (async from pinned memory); everything else untouched."""
    non_blocking = device.type == "cuda"
    return {
        key: value.to(device, non_blocking=non_blocking) if torch.is_tensor(value) else value
        for key, value in batch.items()
    }


class TorchEngine(TrainEngine):
    def __init__(
        self,
        model: nn.Module,
        adapter: Any,
        *,
        device: torch.device,
        lr: float,
        weight_decay: float,
        max_grad_norm: float,
        data_parallel: bool,
        find_unused_parameters: bool = False,
    ) -> None:
        self._adapter = adapter
        self._device = device
        self._max_grad_norm = max_grad_norm
        if hasattr(model, "prefer_native"):
            model.prefer_native = False  # this engine differentiates through torch ops
        self.model = model
        if data_parallel:
            ids = [device.index] if device.type == "cuda" and device.index is not None else None
            self.model = DistributedDataParallel(
                model, device_ids=ids, find_unused_parameters=find_unused_parameters
            )
        self.optimizer = torch.optim.AdamW(
            self.model.parameters(), lr=lr, weight_decay=weight_decay
        )

    @property
    def raw_model(self) -> nn.Module:
        if isinstance(self.model, DistributedDataParallel):
            return self.model.module
        return self.model

    def train_step(self, next_batch: Callable[[], Batch], accum_steps: int) -> tuple[Any, int]:
        self.optimizer.zero_grad()
        loss_total = 0.0
        tokens = 0
        wrapped = isinstance(self.model, DistributedDataParallel)
        for micro in range(accum_steps):
            batch = move_batch(next_batch(), self._device)
            tokens += batch["input_ids"].numel()
            defer_sync = wrapped and micro != accum_steps - 1
            with self.model.no_sync() if defer_sync else contextlib.nullcontext():
                loss, metrics = self._adapter.compute_loss(self.model, batch)
                (loss / accum_steps).backward()
            loss_total += float(metrics.get("loss", loss.item()))
        torch.nn.utils.clip_grad_norm_(self.model.parameters(), self._max_grad_norm)
        self.optimizer.step()
        return loss_total / accum_steps, tokens

    def eval_loss(self, batch: Batch) -> float:
        loss, _ = self._adapter.compute_loss(self.raw_model, move_batch(batch, self._device))
        return float(loss.item())

    def peak_memory_bytes(self) -> float:
        if self._device.type == "cuda":
            return float(torch.cuda.max_memory_allocated(self._device))
        return 0.0
