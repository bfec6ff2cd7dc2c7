"""Autoregressive sampling from a trained :class:`~llmtrain.models.gpt.GPT`.

The reference ships this capability only as a notebook
(`notebooks/trained_vs_random_completion.ipynb`: load ``step_NNNNNN.pt``, then a greedy/top-k loop
that re-runs the full prefix each step, no KV cache).  Here it is a small library function with
the same knobs — ``temperature``, ``top_k`` — plus a checkpoint loader, usable on CPU and on CUDA
(where each forward pass runs on the native kernels).
"""

from __future__ import annotations

from pathlib import Path
from typing import Any

import torch

from llmtrain.config.schemas import RunConfig
from llmtrain.models.gpt import GPT, GPTAdapter


def load_model_from_checkpoint(
    path: str | Path, *, device: str | torch.device = "cpu"
) -> tuple[GPT, RunConfig]:
    """Rebuild the GPT described by a checkpoint's stored config and load its weights."""
    payload: dict[str, Any] = torch.load(Path(path), map_location="cpu", weights_only=False)
    cfg = RunConfig.model_validate(payload["config"])
    if cfg.model.name != "gpt":
        raise ValueError(f"checkpoint was written by model adapter {cfg.model.name!r}, not 'gpt'")
    vocab = payload["model_state_dict"]["token_embedding.weight"].shape[0]
    cfg = cfg.model_copy(update={"model": cfg.model.model_copy(update={"vocab_size": vocab})})
    model = GPTAdapter().build_model(cfg)
    model.load_state_dict(payload["model_state_dict"])
    return model.to(device).eval(), cfg  # type: ignore[return-value]


@torch.no_grad()
def generate(
    model: GPT,
    input_ids: torch.Tensor,
    *,
    max_new_tokens: int,
    temperature: float = 1.0,
    top_k: int | None = None,
    generator: torch.Generator | None = None,
) -> torch.Tensor:
    """Append ``max_new_tokens`` sampled tokens to ``input_ids`` (shape ``(B, T0)``).

    ``temperature == 0`` means greedy decoding.  The context is cropped to the model's
    ``block_size`` (learned absolute positions cannot extrapolate).
    """
    if input_ids.dim() != 2:
        raise ValueError("input_ids must be (batch, time)")
    was_training = model.training
    model.eval()
    tokens = input_ids
    try:
        for _ in range(max_new_tokens):
            window = tokens[:, -model.block_size :]
            logits = model(window)[:, -1, :].float()
            if temperature <= 0.0:
                nxt = logits.argmax(dim=-1, keepdim=True)
            else:
                logits = logits / temperature
                if top_k is not None and 0 < top_k < logits.size(-1):
                    kth = logits.topk(top_k, dim=-1).values[:, -1:]
                    logits = logits.masked_fill(logits < kth, float("-inf"))
                probs = logits.softmax(dim=-1)
                nxt = torch.multinomial(probs, num_samples=1, generator=generator)
            tokens = torch.cat([tokens, nxt.to(tokens.dtype)], dim=1)
    finally:
        model.train(was_training)
    return tokens
