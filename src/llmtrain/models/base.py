"""Model plugin contract (reference: `src/llmtrain/models/base.py:12-27`)."""

from __future__ import annotations

import abc
from typing import Any

import torch
from torch import nn

from llmtrain.config.schemas import RunConfig


class ModelAdapter(abc.ABC):
    """Builds a model + tokenizer from a config and knows how to score one batch."""

    @abc.abstractmethod
    def build_model(self, cfg: RunConfig) -> nn.Module:
        """Construct the (uninitialised-device) model for ``cfg``."""

    @abc.abstractmethod
    def build_tokenizer(self, cfg: RunConfig) -> Any | None:
        """Tokenizer for ``cfg``; ``None`` when the model needs none."""

    @abc.abstractmethod
    def compute_loss(
        self, model: nn.Module, batch: dict[str, torch.Tensor]
    ) -> tuple[torch.Tensor, dict[str, float]]:
        """Scalar loss tensor (differentiable) plus float metrics, at least ``"loss"``."""
