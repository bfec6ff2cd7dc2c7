"""Learning-rate schedule: linear warm-up then cosine decay to zero.

Same curve as the reference (`src/llmtrain/training/trainer.py:104-121`): the multiplier is
``step / warmup`` below ``warmup`` (so LR is 0 at step 0 when warm-up is on), cosine from 1 to 0
between ``warmup`` and ``max_steps``, and exactly 0 from ``max_steps`` on.
"""

from __future__ import annotations

import math
from collections.abc import Callable

import torch


def warmup_cosine_multiplier(warmup_steps: int, max_steps: int) -> Callable[[int], float]:
    """Return ``f(step) -> multiplier`` as a plain function.

    A *function* (not a callable object) on purpose: ``LambdaLR.state_dict()`` then stores
    ``lr_lambdas: [None]``, which is the checkpoint layout the reference writes and reads.
    """
    decay_span = max_steps - warmup_steps

    def multiplier(step: int) -> float:
        if step < warmup_steps:
            return step / warmup_steps if warmup_steps > 0 else 1.0
        if step >= max_steps:
            return 0.0
        if decay_span <= 0:
            return 1.0
        return 0.5 * (1.0 + math.cos(math.pi * (step - warmup_steps) / decay_span))

    return multiplier


def build_scheduler(
    optimizer: torch.optim.Optimizer, warmup_steps: int, max_steps: int
) -> torch.optim.lr_scheduler.LambdaLR:
    """``LambdaLR`` over :func:`warmup_cosine_multiplier`."""
    return torch.optim.lr_scheduler.LambdaLR(
        optimizer, lr_lambda=warmup_cosine_multiplier(warmup_steps, max_steps)
    )
