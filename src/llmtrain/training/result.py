"""Outcome record of a ``Trainer.fit`` call (reference: `src/llmtrain/training/trainer.py:30-43`)."""

from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class TrainResult:
    """What a training run reports back to the CLI / caller."""

    final_step: int
    final_loss: float
    final_val_loss: float | None
    total_time: float
    peak_memory: float
    val_metrics: dict[str, float] | None = None
    first_step_loss: float | None = None
    resumed_from_step: int | None = None
    parameter_count: int | None = None
    trainable_parameter_count: int | None = None
