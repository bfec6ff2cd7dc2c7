"""``dummy_gpt`` — a deliberately tiny, non-causal fixture model.

Plumbing fixture, not a performance target: it stays on stock torch modules
on every device.  Behaviour kept from the reference (`src/llmtrain/models/dummy_gpt.py:13-90`):
vocab defaults to 128, width is capped at 64, one ``nn.TransformerEncoderLayer``, untied head,
and the loss is a plain mean cross-entropy that ignores the mask.
"""

from __future__ import annotations

from typing import Any

import torch
from torch import nn
from torch.nn import functional as F

from llmtrain.config.schemas import RunConfig
from llmtrain.models._checks import validate_lm_batch
from llmtrain.models.base import ModelAdapter
from llmtrain.registry.models import register_model

_WIDTH_CAP = 64


class _TinyGPT(nn.Module):
    def __init__(self, vocab_size: int, d_model: int, n_heads: int) -> None:
        super().__init__()
        self.embed = nn.Embedding(vocab_size, d_model)
        layer = nn.TransformerEncoderLayer(d_model=d_model, nhead=n_heads, batch_first=True)
        self.encoder = nn.TransformerEncoder(layer, num_layers=1, enable_nested_tensor=False)
        self.lm_head = nn.Linear(d_model, vocab_size, bias=False)

    def forward(
        self, input_ids: torch.Tensor, attention_mask: torch.Tensor | None = None
    ) -> torch.Tensor:
        pad = None if attention_mask is None else attention_mask == 0
        return self.lm_head(self.encoder(self.embed(input_ids), src_key_padding_mask=pad))


@register_model("dummy_gpt")
class DummyGPTAdapter(ModelAdapter):
    """Adapter for the fixture model."""

    def build_model(self, cfg: RunConfig) -> nn.Module:
        width = min(cfg.model.d_model or 128, _WIDTH_CAP)
        heads = max(1, min(cfg.model.n_heads, width))
        if width % heads:
            heads = 1 if width % 2 else 2
        return _TinyGPT(vocab_size=cfg.model.vocab_size or 128, d_model=width, n_heads=heads)

    def build_tokenizer(self, cfg: RunConfig) -> Any | None:
        del cfg
        return None

    def compute_loss(
        self, model: nn.Module, batch: dict[str, torch.Tensor]
    ) -> tuple[torch.Tensor, dict[str, float]]:
        ids, labels, mask = validate_lm_batch(batch, min_seq_len=1)
        logits = model(ids, attention_mask=mask)
        loss = F.cross_entropy(logits.flatten(0, 1), labels.flatten())
        return loss, {"loss": loss.item()}
