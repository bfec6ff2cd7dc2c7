"""``synthetic_text`` — full-length random token rows for throughput work on B200.

The reference cannot produce a sequence longer than 8 tokens offline (its ``dummy_text`` clamps
to 8 / 128 examples — `src/llmtrain/data/dummy_text.py:68-72`; ``hf_text`` downloads).  This
module generates ``T = block_size`` rows with next-token labels (``labels[t] = stream[t+1]``)
and an all-ones mask, i.e. the same row shape ``hf_text`` yields, from a counter-based generator
so that any (epoch, index) row is reproducible without storing the corpus.

Knobs (``data.extra``): ``num_examples`` (default: enough for ``max_steps``, capped at 4096),
``val_examples`` (default 0 → no validation loader), ``pattern`` (``"random"`` or
``"arith"`` — a learnable arithmetic progression used by loss-goes-down tests).

Rows are produced on the host in pinned memory when the run device is CUDA, so the trainer's
host→device copy is asynchronous.  With ``data.extra.device_resident: true`` (CUDA runs, pattern
``random``) the batches are generated ON the device instead, by the library's counter-based
token kernel (``llmtrain_b200_synthetic_tokens``): no host work and no copy at all,
every rank draws its own rows of one global stream.
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


class _SyntheticRows(Dataset[dict[str, torch.Tensor]]):
    def __init__(self, n: int, seq_len: int, vocab: int, seed: int, pattern: str) -> None:
        self._n, self._t, self._v, self._seed, self._pattern = n, seq_len, vocab, seed, pattern

    def __len__(self) -> int:
        return self._n

    def __getitem__(self, index: int) -> dict[str, torch.Tensor]:
        gen = torch.Generator().manual_seed(self._seed * 1_000_003 + index)
        if self._pattern == "arith":
            start = int(torch.randint(0, self._v, (1,), generator=gen))
            stride = 1 + int(torch.randint(0, 3, (1,), generator=gen))
            stream = (start + stride * torch.arange(self._t + 1)) % self._v
        else:
            stream = torch.randint(0, self._v, (self._t + 1,), dtype=torch.long, generator=gen)
        stream = stream.to(torch.long)
        ids = stream[:-1].clone()
        return {"input_ids": ids, "labels": stream[1:].clone(), "attention_mask": torch.ones_like(ids)}


class _DeviceBatches:
    """Endless stream of device-generated ``(input_ids, labels)`` batches (train split only)."""

    def __init__(self, batch: int, seq_len: int, vocab: int, seed: int, rank: int,
                 world: int) -> None:
        self._shape = (batch, seq_len, vocab)
        self._seed, self._rank, self._world = seed, rank, world
        self._batches_per_epoch = 1 << 20

    def __len__(self) -> int:
        return self._batches_per_epoch

    def __iter__(self):  # noqa: ANN204
        from llmtrain.ops import kernels as K

        batch, seq_len, vocab = self._shape
        for index in range(self._batches_per_epoch):
            first_row = (index * self._world + self._rank) * batch
            ids, labels = K.synthetic_tokens(batch, seq_len, vocab, seed=self._seed,
                                             first_row=first_row)
            yield {"input_ids": ids, "labels": labels}


@register_data_module("synthetic_text")
class SyntheticTextDataModule(DataModule):
    """Random (or arithmetic) token rows of length ``block_size``."""

    def __init__(self) -> None:
        self._cfg: RunConfig | None = None
        self._train: _SyntheticRows | None = None
        self._val: _SyntheticRows | None = None

    def setup(self, cfg: RunConfig, tokenizer: Any | None = None) -> None:
        del tokenizer
        extra = cfg.data.extra
        vocab = cfg.model.vocab_size or 50257
        seq_len = int(extra.get("seq_len", cfg.model.block_size))
        if not 2 <= seq_len <= cfg.model.block_size:
            raise ValueError("synthetic_text: data.extra.seq_len must be in [2, model.block_size]")
        per_step = cfg.trainer.micro_batch_size * cfg.trainer.grad_accum_steps
        world = max(1, cfg.ddp.world_size or 1)
        default_n = min(4096, cfg.trainer.max_steps * per_step * world)
        n_train = max(1, int(extra.get("num_examples", default_n)))
        n_val = max(0, int(extra.get("val_examples", 0)))
        pattern = str(extra.get("pattern", "random"))
        if pattern not in ("random", "arith"):
            raise ValueError("synthetic_text: data.extra.pattern must be 'random' or 'arith'")
        self._cfg = cfg
        self._train = _SyntheticRows(n_train, seq_len, vocab, cfg.run.seed, pattern)
        self._val = (
            _SyntheticRows(n_val, seq_len, vocab, cfg.run.seed + 1000, pattern) if n_val else None
        )

    def _loader(self, rows: _SyntheticRows, *, train: bool) -> DataLoader:
        assert self._cfg is not None
        return build_loader(
            rows,
            self._cfg,
            dist,
            train=train,
            num_workers=0,
            pin_memory=self._cfg.run.device == "cuda" and torch.cuda.is_available(),
        )

    def train_dataloader(self) -> DataLoader:
        if self._cfg is None or self._train is None:
            raise RuntimeError("setup must be called before train_dataloader")
        cfg = self._cfg
        if bool(cfg.data.extra.get("device_resident", False)):
            if cfg.run.device != "cuda" or not torch.cuda.is_available():
                raise ValueError("synthetic_text: data.extra.device_resident needs run.device: cuda")
            if self._train._pattern != "random":
                raise ValueError("synthetic_text: device_resident supports pattern 'random' only")
            rank, world = 0, 1
            if dist.is_available() and dist.is_initialized():
                rank, world = dist.get_rank(), dist.get_world_size()
            stream = _DeviceBatches(cfg.trainer.micro_batch_size, self._train._t, self._train._v,
                                    cfg.run.seed, rank, world)
            return stream  # type: ignore[return-value]  # iterable of batches, like a DataLoader
        return self._loader(self._train, train=True)

    def val_dataloader(self) -> DataLoader | None:
        if self._cfg is None:
            raise RuntimeError("setup must be called before val_dataloader")
        return None if self._val is None else self._loader(self._val, train=False)
