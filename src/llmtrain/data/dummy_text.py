"""``dummy_text`` — tiny synthetic token data for plumbing tests.

Semantics kept from the reference (`src/llmtrain/data/dummy_text.py:15-153`): sequences of
``min(block_size, 8)`` tokens, at most 128 training examples, labels *equal* to inputs (so a
model can drive the loss to ~0), all-ones mask, per-index seeding when ``run.deterministic``,
validation set = a fifth of train (1..32) seeded 1000 higher, never any worker processes.
"""

from __future__ import annotations

from typing import Any

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader, Dataset

from llmtrain.config.schemas import RunConfig
from llmtrain.data._sharding import build_loader
from llmtrain.data.base import DataModule
from llmtrain.registry.data import register_data_module

_MAX_SEQ = 8
_MAX_TRAIN = 128
_MAX_VAL = 32
_VAL_SEED_OFFSET = 1000


class _DummyTextDataset(Dataset[dict[str, torch.Tensor]]):
    """Index → random token row; reproducible per index when ``deterministic``."""

    def __init__(
        self, num_examples: int, seq_len: int, vocab_size: int, deterministic: bool, seed: int
    ) -> None:
        self._n = num_examples
        self._t = seq_len
        self._v = vocab_size
        self._deterministic = deterministic
        self._seed = seed

    def __len__(self) -> int:
        return self._n

    def __getitem__(self, index: int) -> dict[str, torch.Tensor]:
        gen: torch.Generator | None = None
        if self._deterministic:
            gen = torch.Generator().manual_seed(self._seed + index)
        ids = torch.randint(0, self._v, (self._t,), dtype=torch.long, generator=gen)
        return {"input_ids": ids, "labels": ids.clone(), "attention_mask": torch.ones_like(ids)}


@register_data_module("dummy_text")
class DummyTextDataModule(DataModule):
    """Synthetic text for dry runs and smoke tests."""

    def __init__(self) -> None:
        self._cfg: RunConfig | None = None
        self._train_dataset: Dataset[dict[str, torch.Tensor]] | None = None
        self._val_dataset: Dataset[dict[str, torch.Tensor]] | None = None

    def setup(self, cfg: RunConfig, tokenizer: Any | None = None) -> None:
        del tokenizer
        vocab = cfg.model.vocab_size or 128
        seq_len = max(1, min(cfg.model.block_size, _MAX_SEQ))
        wanted = (cfg.trainer.max_steps or 1) * (cfg.trainer.micro_batch_size or 1)
        n_train = max(1, min(wanted, _MAX_TRAIN))
        n_val = max(1, min(n_train // 5, _MAX_VAL))
        det, seed = cfg.run.deterministic, cfg.run.seed
        self._cfg = cfg
        self._train_dataset = _DummyTextDataset(n_train, seq_len, vocab, det, seed)
        self._val_dataset = _DummyTextDataset(n_val, seq_len, vocab, det, seed + _VAL_SEED_OFFSET)

    def train_dataloader(self) -> DataLoader:
        if self._cfg is None or self._train_dataset is None:
            raise RuntimeError("setup must be called before train_dataloader")
        return build_loader(self._train_dataset, self._cfg, dist, train=True, num_workers=0)

    def val_dataloader(self) -> DataLoader | None:
        if self._cfg is None:
            raise RuntimeError("setup must be called before val_dataloader")
        if self._val_dataset is None:
            return None
        return build_loader(self._val_dataset, self._cfg, dist, train=False, num_workers=0)
