"""Rank-sharding policy shared by every data module.

A ``DistributedSampler`` is used when a real process group with more than one rank is up,
*or* when only the config hints ``ddp.world_size > 1`` (lets tests and offline tooling build
rank-specific loaders without a process group).  Behavioural parity with the reference
(`src/llmtrain/data/dummy_text.py:96-113`, `src/llmtrain/data/hf_text.py:181-198`).

The owning module must expose ``torch.distributed`` as its own ``dist`` attribute and pass it
in, so tests can patch ``llmtrain.data.<module>.dist``.
"""

from __future__ import annotations

from typing import Any

from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler

from llmtrain.config.schemas import RunConfig


def build_loader(
    dataset: Any,
    cfg: RunConfig,
    dist_module: Any,
    *,
    train: bool,
    num_workers: int,
    collate_fn: Any | None = None,
    pin_memory: bool = False,
) -> DataLoader:
    """Create the loader for one split with the reference's shuffle/sampler rules."""
    world = cfg.ddp.world_size or 1
    rank = cfg.ddp.rank or 0
    if dist_module.is_available() and dist_module.is_initialized():
        world, rank = dist_module.get_world_size(), dist_module.get_rank()
    want_shuffle = train and not cfg.run.deterministic

    sampler: DistributedSampler | None = None
    if world > 1:
        sampler = DistributedSampler(
            dataset,
            num_replicas=world,
            rank=rank,
            shuffle=want_shuffle,
            seed=cfg.run.seed,
        )
    kwargs: dict[str, Any] = {}
    if collate_fn is not None:
        kwargs["collate_fn"] = collate_fn
    return DataLoader(
        dataset,
        batch_size=cfg.trainer.micro_batch_size or 1,
        num_workers=num_workers,
        shuffle=want_shuffle and sampler is None,
        sampler=sampler,
        pin_memory=pin_memory,
        **kwargs,
    )
