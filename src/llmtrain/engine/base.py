"""Execution-engine interface the :class:`~llmtrain.training.trainer.Trainer` drives.

The reference hard-wires its optimisation step to autograd + ``DistributedDataParallel`` +
``torch.optim.AdamW`` (`src/llmtrain/training/trainer.py:361-395`).  Here that step is a
strategy object with two implementations:

* :class:`llmtrain.engine.torch_engine.TorchEngine` — the same autograd/DDP/AdamW sequence on
  torch ops (CPU, MPS, or CUDA with ``ddp.backend`` gloo/nccl); the oracle and plumbing path.
* :class:`llmtrain.engine.b200.B200TrainEngine` — the sm_100a runtime: C++ GPT engine, bf16
  gradient arena, NVLink peer-memory reduce fused with clip + AdamW.
"""

from __future__ import annotations

import abc
from collections.abc import Callable
from typing import Any

import torch
from torch import nn

Batch = dict[str, Any]


class TrainEngine(abc.ABC):
    """One optimiser step at a time."""

    #: what the trainer exposes as ``_model`` (may be a DDP wrapper)
    model: nn.Module
    #: state container whose ``state_dict()`` goes into checkpoints
    optimizer: torch.optim.Optimizer

    @property
    @abc.abstractmethod
    def raw_model(self) -> nn.Module:
        """The user's module without any wrapper."""

    @abc.abstractmethod
    def train_step(self, next_batch: Callable[[], Batch], accum_steps: int) -> tuple[Any, int]:
        """Run ``accum_steps`` micro-batches + one optimiser update.

        Returns ``(mean micro-batch loss, tokens consumed on this rank)``.  The loss may be a
        float, a 0-dim device tensor, or any object with ``__float__`` whose value arrives later
        (the B200 engine returns ``StepScalars``: a pinned-host copy guarded by an event); the
        trainer calls ``float()`` on it only when a value is needed on the host (first step, log
        lines — written one step late —, last step).
        """

    @abc.abstractmethod
    def eval_loss(self, batch: Batch) -> float:
        """Forward-only loss of one batch (model already in eval mode)."""

    def before_checkpoint(self) -> None:  # noqa: B027 - optional hook
        """Make ``model.state_dict()`` / ``optimizer.state_dict()`` current."""

    def after_restore(self) -> None:  # noqa: B027 - optional hook
        """Parameters / optimiser state were overwritten in place by a checkpoint load."""

    def peak_memory_bytes(self) -> float:
        return 0.0

    def extra_metrics(self, tokens_per_sec: float | None = None,
                      step_result: Any = None) -> dict[str, float]:
        """Engine-specific metrics appended (never renaming the reference's keys).

        ``tokens_per_sec`` is this rank's rate over the interval being logged, for engines that
        report utilisation; ``step_result`` is what ``train_step`` returned for the interval's
        last step (log lines are written one step late, see ``Trainer.fit``).
        """
        del tokens_per_sec, step_result
        return {}
