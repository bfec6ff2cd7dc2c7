"""``hf_text`` — Hugging Face text dataset → fixed-length next-token training rows.

Pipeline and cache layout follow the reference (`src/llmtrain/data/hf_text.py:36-240`):
tokenise every row, concatenate inside each ``map`` batch, cut into ``block_size + 1`` windows,
emit ``input_ids = w[:-1]``, ``labels = w[1:]`` (pre-shifted), all-ones mask; the processed
split is stored under ``cache_dir/processed/{name}__{config}__b{block}__{split}`` and re-used.
"""

from __future__ import annotations

from pathlib import Path
from typing import Any

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader

from llmtrain.config.schemas import RunConfig
from llmtrain.data._sharding import build_loader
from llmtrain.data.base import DataModule
from llmtrain.registry.data import register_data_module

_COLUMNS = ("input_ids", "labels", "attention_mask")


def _encode_rows(tokenizer: Any, texts: list[Any]) -> list[list[int]]:
    rows: list[list[int]] = []
    for text in texts:
        if text is None:
            rows.append([])
            continue
        ids = tokenizer.encode(str(text))
        if not isinstance(ids, list):
            raise ValueError("Tokenizer encode output must be a list of token ids.")
        rows.append([int(tok) for tok in ids])
    return rows


def _window(token_rows: list[list[int]], block_size: int) -> dict[str, list[list[int]]]:
    stream = [tok for row in token_rows for tok in row]
    width = block_size + 1
    n_windows = len(stream) // width
    out: dict[str, list[list[int]]] = {name: [] for name in _COLUMNS}
    ones = [1] * block_size
    for w in range(n_windows):
        piece = stream[w * width : (w + 1) * width]
        out["input_ids"].append(piece[:-1])
        out["labels"].append(piece[1:])
        out["attention_mask"].append(list(ones))
    return out


@register_data_module("hf_text")
class HFTextDataModule(DataModule):
    """Tokenised + chunked Hugging Face dataset with an on-disk processed cache."""

    def __init__(self) -> None:
        self._cfg: RunConfig | None = None
        self._train_dataset: Any | None = None
        self._val_dataset: Any | None = None

    @staticmethod
    def _collate_batch(batch: list[dict[str, list[int]]]) -> dict[str, torch.Tensor]:
        return {
            name: torch.tensor([row[name] for row in batch], dtype=torch.long) for name in _COLUMNS
        }

    def setup(self, cfg: RunConfig, tokenizer: Any | None = None) -> None:
        if cfg.data.dataset_name is None:
            raise ValueError("hf_text requires data.dataset_name to be configured.")
        if cfg.data.text_column is None:
            raise ValueError("hf_text requires data.text_column to be configured.")
        if tokenizer is None:
            raise ValueError("hf_text requires a tokenizer instance.")
        if not hasattr(tokenizer, "encode"):
            raise ValueError("hf_text tokenizer must provide an encode(text) method.")

        import datasets  # type: ignore[import-untyped]  # resolved late so tests can patch it

        self._cfg = cfg
        splits = {}
        for split in (cfg.data.train_split, cfg.data.val_split):
            splits[split] = self._prepare_split(split, cfg, tokenizer, datasets)
        self._train_dataset = splits[cfg.data.train_split]
        self._val_dataset = splits[cfg.data.val_split]

    def _processed_cache_path(self, cfg: RunConfig, split: str) -> Path:
        name = (cfg.data.dataset_name or "unknown").replace("/", "__")
        variant = (cfg.data.dataset_config or "default").replace("/", "__")
        leaf = f"{name}__{variant}__b{cfg.model.block_size}__{split}"
        return Path(cfg.data.cache_dir) / "processed" / leaf

    def _prepare_split(self, split: str, cfg: RunConfig, tokenizer: Any, hub: Any) -> Any:
        cached = self._processed_cache_path(cfg, split)
        if cached.exists():
            return hub.load_from_disk(str(cached))
        raw = hub.load_dataset(
            cfg.data.dataset_name,
            cfg.data.dataset_config,
            split=split,
            cache_dir=cfg.data.cache_dir,
        )
        processed = self._tokenize_and_chunk(
            raw_dataset=raw,
            tokenizer=tokenizer,
            text_column=str(cfg.data.text_column),
            block_size=cfg.model.block_size,
        )
        cached.parent.mkdir(parents=True, exist_ok=True)
        processed.save_to_disk(str(cached))
        return processed

    def _tokenize_and_chunk(
        self, *, raw_dataset: Any, tokenizer: Any, text_column: str, block_size: int
    ) -> Any:
        tokenised = raw_dataset.map(
            lambda rows: {"token_ids": _encode_rows(tokenizer, rows[text_column])},
            batched=True,
            remove_columns=raw_dataset.column_names,
            desc="Tokenizing",
        )
        return tokenised.map(
            lambda rows: _window(rows["token_ids"], block_size),
            batched=True,
            remove_columns=["token_ids"],
            desc="Chunking",
        )

    def _loader(self, dataset: Any, *, train: bool) -> DataLoader:
        assert self._cfg is not None
        return build_loader(
            dataset,
            self._cfg,
            dist,
            train=train,
            num_workers=self._cfg.data.num_workers,
            collate_fn=self._collate_batch,
            pin_memory=self._cfg.run.device == "cuda",
        )

    def train_dataloader(self) -> DataLoader:
        if self._cfg is None or self._train_dataset is None:
            raise RuntimeError("setup must be called before train_dataloader")
        return self._loader(self._train_dataset, train=True)

    def val_dataloader(self) -> DataLoader | None:
        if self._cfg is None:
            raise RuntimeError("setup must be called before val_dataloader")
        if self._val_dataset is None:
            return None
        return self._loader(self._val_dataset, train=False)
