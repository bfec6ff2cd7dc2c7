"""Batch validation shared by the adapters.

Error texts match the reference adapters (`src/llmtrain/models/gpt.py:221-252`,
`src/llmtrain/models/dummy_gpt.py:59-87`) because callers and tests match on them.
"""

from __future__ import annotations

import torch


def validate_lm_batch(
    batch: dict[str, torch.Tensor], *, min_seq_len: int
) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor | None]:
    """Return ``(input_ids, labels, attention_mask)`` after shape/dtype checks."""
    ids = batch["input_ids"]
    labels = batch["labels"]
    mask = batch.get("attention_mask")

    if ids.dim() != 2 or labels.dim() != 2:
        raise ValueError(
            "Expected input_ids and labels to be 2D (B, T); "
            f"got {tuple(ids.shape)} and {tuple(labels.shape)}."
        )
    if ids.shape != labels.shape:
        raise ValueError(
            "Expected input_ids and labels to have the same shape; "
            f"got {tuple(ids.shape)} vs {tuple(labels.shape)}."
        )
    if ids.dtype != torch.long or labels.dtype != torch.long:
        raise ValueError(
            f"Expected input_ids and labels to be torch.long; got {ids.dtype} and {labels.dtype}."
        )
    if ids.size(1) < min_seq_len:
        raise ValueError(f"Expected sequence length >= {min_seq_len} for next-token loss.")
    if mask is not None:
        if mask.dim() != 2:
            raise ValueError(f"Expected attention_mask to be 2D (B, T); got {tuple(mask.shape)}.")
        if mask.shape != ids.shape:
            raise ValueError(
                "Expected attention_mask to match input_ids shape; "
                f"got {tuple(mask.shape)} vs {tuple(ids.shape)}."
            )
        if mask.dtype not in (torch.bool, torch.int64):
            raise ValueError(f"Expected attention_mask to be bool or int64; got {mask.dtype}.")
    return ids, labels, mask
