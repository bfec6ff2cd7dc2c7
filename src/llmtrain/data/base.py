"""Data plugin contract (reference: `src/llmtrain/data/base.py:11-24`)."""

from __future__ import annotations

import abc
from typing import Any

from torch.utils.data import DataLoader

from llmtrain.config.schemas import RunConfig


class DataModule(abc.ABC):
    """Owns datasets; hands out loaders that yield ``input_ids/labels/attention_mask``."""

    @abc.abstractmethod
    def setup(self, cfg: RunConfig, tokenizer: Any | None = None) -> None:
        """Materialise datasets for ``cfg``."""

    @abc.abstractmethod
    def train_dataloader(self) -> DataLoader:
        """Training loader (rank-sharded when data-parallel)."""

    @abc.abstractmethod
    def val_dataloader(self) -> DataLoader | None:
        """Validation loader or ``None``."""
