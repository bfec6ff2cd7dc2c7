"""Decoder-only GPT: parameter container + two execution paths.

*What* is computed is fixed by the reference (`src/llmtrain/models/gpt.py:15-271`; spelled out
in SURVEY §2.4): learned absolute positions, pre-LN blocks, fused QKV projection, causal
softmax attention with ``finfo.min`` masking, exact-erf GELU MLP, final LayerNorm, (tied) LM
head, N(0, 0.02²) init with residual projections scaled by ``1/sqrt(2L)``.  The module tree and
``state_dict`` keys are identical, including the per-layer ``attn.causal_mask`` entry, so
checkpoints interchange with the reference in both directions.

*How* it is computed depends on where the parameters live:

* CPU (or MPS): plain torch ops, fp32 — this is the numerical oracle and the plumbing path.
* CUDA (B200, sm_100a): the module hands its parameters to :class:`llmtrain.engine.b200.GPTEngine`,
  a C++ runtime that owns flat parameter/gradient/activation arenas and launches this repo's
  hand-written tcgen05/TMA kernels (``csrc/``).  ``forward`` (under ``no_grad``) returns logits
  produced by those kernels and ``GPTAdapter.compute_loss`` runs the fused forward+backward.
  A missing native library is an error, never a silent fallback.  The torch-op path remains
  reachable on CUDA on purpose (``prefer_native = False`` / ``trainer.extra.engine: torch``):
  it is the on-device oracle for the equivalence tests and the path taken for shapes the kernels
  do not cover (head_dim > 64).

The causal mask is never materialised on the device (the attention kernel applies it
implicitly); the ``causal_mask`` tensors exist only inside ``state_dict()`` for format parity,
and all layers share one storage so a checkpoint stores it once.
"""

from __future__ import annotations

import math
from functools import lru_cache
from typing import Any

import torch
from torch import nn
from torch.nn import functional as F

from llmtrain.config.schemas import RunConfig
from llmtrain.models._checks import validate_lm_batch
from llmtrain.models.base import ModelAdapter
from llmtrain.models.tokenizer import load_gpt2_tokenizer
from llmtrain.registry.models import register_model

_INIT_STD = 0.02


@lru_cache(maxsize=8)
def _strict_upper_mask(block_size: int) -> torch.Tensor:
    """(1,1,T,T) bool, True above the diagonal — host tensor shared by every layer."""
    return torch.ones(block_size, block_size, dtype=torch.bool).triu_(1)[None, None]


class CausalSelfAttention(nn.Module):
    """Multi-head causal self-attention (oracle math; fused kernel on CUDA)."""

    def __init__(self, d_model: int, n_heads: int, block_size: int, dropout: float) -> None:
        super().__init__()
        if d_model % n_heads:
            raise ValueError("d_model must be divisible by n_heads")
        self.d_model = d_model
        self.n_heads = n_heads
        self.head_dim = d_model // n_heads
        self.block_size = block_size
        self.qkv_proj = nn.Linear(d_model, 3 * d_model)
        self.attn_dropout = nn.Dropout(dropout)
        self.out_proj = nn.Linear(d_model, d_model)
        self.resid_dropout = nn.Dropout(dropout)

    # -- checkpoint-format parity: a virtual persistent buffer ---------------------------------
    @property
    def causal_mask(self) -> torch.Tensor:
        return _strict_upper_mask(self.block_size)

    def _save_to_state_dict(self, destination: Any, prefix: str, keep_vars: bool) -> None:
        super()._save_to_state_dict(destination, prefix, keep_vars)
        destination[prefix + "causal_mask"] = self.causal_mask

    def _load_from_state_dict(  # type: ignore[no-untyped-def]
        self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs
    ) -> None:
        super()._load_from_state_dict(
            state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs
        )
        key = prefix + "causal_mask"
        if key in unexpected_keys:
            unexpected_keys.remove(key)
        stored = state_dict.get(key)
        if stored is not None and tuple(stored.shape[-2:]) != (self.block_size, self.block_size):
            error_msgs.append(
                f"size mismatch for {key}: checkpoint block size {tuple(stored.shape[-2:])} "
                f"vs model block size {self.block_size}"
            )

    # -- oracle path ---------------------------------------------------------------------------
    def forward(self, x: torch.Tensor, attention_mask: torch.Tensor | None = None) -> torch.Tensor:
        batch, seq, _ = x.shape
        if seq > self.block_size:
            raise ValueError(f"Input sequence length {seq} exceeds block size {self.block_size}.")
        if attention_mask is not None and tuple(attention_mask.shape) != (batch, seq):
            raise ValueError(
                "Expected attention_mask to have shape (B, T); "
                f"got {tuple(attention_mask.shape)} for {(batch, seq)}."
            )
        heads, hd = self.n_heads, self.head_dim
        # (B,T,3,H,hd) -> three (B,H,T,hd) views
        q, k, v = self.qkv_proj(x).view(batch, seq, 3, heads, hd).permute(2, 0, 3, 1, 4)

        scores = torch.matmul(q, k.transpose(-1, -2)) * (1.0 / math.sqrt(hd))
        lowest = torch.finfo(scores.dtype).min  # not -inf: a fully masked row stays finite
        future = torch.ones(seq, seq, dtype=torch.bool, device=x.device).triu_(1)
        scores = scores.masked_fill(future, lowest)
        keep: torch.Tensor | None = None
        if attention_mask is not None:
            keep = attention_mask.bool()
            scores = scores.masked_fill(~keep[:, None, None, :], lowest)

        probs = self.attn_dropout(scores.softmax(dim=-1))
        mixed = torch.matmul(probs, v).transpose(1, 2).reshape(batch, seq, self.d_model)
        out = self.resid_dropout(self.out_proj(mixed))
        if keep is not None:
            out = out * keep[:, :, None].to(out.dtype)
        return out


class TransformerBlock(nn.Module):
    """Pre-LayerNorm residual block: attention then GELU MLP."""

    def __init__(
        self, d_model: int, n_heads: int, d_ff: int, block_size: int, dropout: float
    ) -> None:
        super().__init__()
        self.ln_1 = nn.LayerNorm(d_model)
        self.attn = CausalSelfAttention(d_model, n_heads, block_size, dropout)
        self.ln_2 = nn.LayerNorm(d_model)
        self.mlp_fc = nn.Linear(d_model, d_ff)
        self.mlp_act = nn.GELU()
        self.mlp_proj = nn.Linear(d_ff, d_model)
        self.mlp_dropout = nn.Dropout(dropout)

    def forward(self, x: torch.Tensor, attention_mask: torch.Tensor | None = None) -> torch.Tensor:
        x = x + self.attn(self.ln_1(x), attention_mask=attention_mask)
        hidden = self.mlp_act(self.mlp_fc(self.ln_2(x)))
        return x + self.mlp_dropout(self.mlp_proj(hidden))


class GPT(nn.Module):
    """Decoder-only language model."""

    def __init__(
        self,
        vocab_size: int,
        block_size: int,
        d_model: int,
        n_layers: int,
        n_heads: int,
        d_ff: int,
        dropout: float,
        tie_embeddings: bool = True,
    ) -> None:
        super().__init__()
        self.vocab_size = vocab_size
        self.block_size = block_size
        self.d_model = d_model
        self.n_layers = n_layers
        self.n_heads = n_heads
        self.d_ff = d_ff
        self.dropout_p = dropout
        self.tie_embeddings = tie_embeddings

        self.token_embedding = nn.Embedding(vocab_size, d_model)
        self.position_embedding = nn.Embedding(block_size, d_model)
        self.drop = nn.Dropout(dropout)
        self.blocks = nn.ModuleList(
            TransformerBlock(d_model, n_heads, d_ff, block_size, dropout) for _ in range(n_layers)
        )
        self.ln_f = nn.LayerNorm(d_model)
        self.lm_head = nn.Linear(d_model, vocab_size, bias=False)
        if tie_embeddings:
            self.lm_head.weight = self.token_embedding.weight
        self.reset_parameters()
        self._engine: Any | None = None  # created lazily once parameters sit on a CUDA device
        #: engines that want the torch-op (autograd) path on CUDA, e.g. the oracle, clear this
        self.prefer_native = True

    def reset_parameters(self) -> None:
        """GPT-2 style init; residual-path projections get the depth-scaled std."""
        for module in self.modules():
            if isinstance(module, (nn.Linear, nn.Embedding)):
                nn.init.normal_(module.weight, mean=0.0, std=_INIT_STD)
                if isinstance(module, nn.Linear) and module.bias is not None:
                    nn.init.zeros_(module.bias)
        depth_std = _INIT_STD / math.sqrt(2 * self.n_layers)
        for block in self.blocks:
            nn.init.normal_(block.attn.out_proj.weight, mean=0.0, std=depth_std)
            nn.init.normal_(block.mlp_proj.weight, mean=0.0, std=depth_std)

    # -- native engine plumbing ---------------------------------------------------------------
    @property
    def on_cuda(self) -> bool:
        return self.token_embedding.weight.is_cuda

    def use_native(self) -> bool:
        """True when calls should run on the sm_100a runtime instead of torch ops.

        Dropout is native too (counter-based masks regenerated in backward, see
        ``csrc/common/philox.cuh``), so training mode with ``dropout > 0`` stays on this path.
        """
        if not (self.on_cuda and self.prefer_native):
            return False
        from llmtrain.engine.b200 import supports

        return supports(self)

    def native_engine(self, **options: Any) -> Any:
        """The C++/CUDA runtime bound to this module's parameters (CUDA only)."""
        if not self.on_cuda:
            raise RuntimeError("native_engine() needs the model on a CUDA device")
        if self._engine is None or not self._engine.bound_to(self):
            from llmtrain.engine.b200 import GPTEngine

            self._engine = GPTEngine(self, **options)
        return self._engine

    def _apply(self, fn: Any, recurse: bool = True) -> GPT:  # type: ignore[override]
        # .to()/.cuda()/.cpu() re-home the parameters: any arena views are stale afterwards.
        self._engine = None
        return super()._apply(fn, recurse)

    # -- forward ---------------------------------------------------------------------------------
    def forward(
        self, input_ids: torch.Tensor, attention_mask: torch.Tensor | None = None
    ) -> torch.Tensor:
        _, seq = input_ids.shape
        if seq > self.block_size:
            raise ValueError(f"Input sequence length {seq} exceeds block size {self.block_size}.")
        if self.use_native() and not torch.is_grad_enabled():
            return self.native_engine().logits(input_ids, attention_mask)
        positions = torch.arange(seq, device=input_ids.device)
        x = self.drop(self.token_embedding(input_ids) + self.position_embedding(positions)[None])
        for block in self.blocks:
            x = block(x, attention_mask=attention_mask)
        return self.lm_head(self.ln_f(x))


@register_model("gpt")
class GPTAdapter(ModelAdapter):
    """Adapter for :class:`GPT`."""

    def build_model(self, cfg: RunConfig) -> nn.Module:
        m = cfg.model
        vocab = m.vocab_size
        if vocab is None:
            vocab = getattr(self.build_tokenizer(cfg), "n_vocab", None)
            if not isinstance(vocab, int) or vocab <= 0:
                raise ValueError("GPT tokenizer must expose a positive integer n_vocab.")
        return GPT(
            vocab_size=vocab,
            block_size=m.block_size,
            d_model=m.d_model,
            n_layers=m.n_layers,
            n_heads=m.n_heads,
            d_ff=m.d_ff,
            dropout=m.dropout,
            tie_embeddings=m.tie_embeddings,
        )

    def build_tokenizer(self, cfg: RunConfig) -> Any | None:
        del cfg
        return load_gpt2_tokenizer()

    def compute_loss(
        self, model: nn.Module, batch: dict[str, torch.Tensor]
    ) -> tuple[torch.Tensor, dict[str, float]]:
        ids, labels, mask = validate_lm_batch(batch, min_seq_len=2)
        inner = getattr(model, "module", model)  # tolerate a DDP wrapper
        if isinstance(inner, GPT) and inner.use_native():
            # Fused native path: bf16 logits live only in the engine's activation arena (consumed in
            # place by the loss kernel); loss keeps an autograd edge to params.
            loss = inner.native_engine().loss(ids, labels, mask)
            return loss, {"loss": float(loss.item())}

        logits = model(ids, attention_mask=mask)
        per_token = F.cross_entropy(logits.flatten(0, 1), labels.flatten(), reduction="none")
        if mask is None:
            loss = per_token.mean()
        else:
            valid = mask.bool().flatten()
            if int(valid.sum().item()) == 0:
                raise ValueError("attention_mask has no valid target tokens.")
            loss = per_token[valid].mean()
        return loss, {"loss": float(loss.item())}
