"""Offline-safe GPT-2 tokenizer access.

The reference calls ``tiktoken.get_encoding("gpt2")`` unconditionally
(`src/llmtrain/models/gpt.py:210-212`), which needs the network the first time and therefore
kills runs that never tokenise anything (``dummy_text`` / ``synthetic_text``).  Here the real
encoding is used when it can be loaded; otherwise a stand-in is returned that knows the GPT-2
vocabulary size and fails only if somebody actually tries to encode or decode with it.
"""

from __future__ import annotations

import logging
from typing import Any

GPT2_VOCAB_SIZE = 50257

logger = logging.getLogger(__name__)


class UnavailableTokenizer:
    """Placeholder with the right ``n_vocab`` that raises on first real use."""

    n_vocab = GPT2_VOCAB_SIZE
    name = "gpt2-unavailable"

    def __init__(self, reason: str) -> None:
        self._reason = reason

    def _fail(self) -> Any:
        raise RuntimeError(
            "The GPT-2 BPE tables could not be loaded (offline and not cached?): "
            f"{self._reason}. Set TIKTOKEN_CACHE_DIR to a directory that holds them."
        )

    def encode(self, text: str, **kwargs: Any) -> list[int]:
        del text, kwargs
        return self._fail()

    def decode(self, tokens: list[int], **kwargs: Any) -> str:
        del tokens, kwargs
        return self._fail()


_cached: Any | None = None


def load_gpt2_tokenizer() -> Any:
    """Return tiktoken's ``gpt2`` encoding, or an :class:`UnavailableTokenizer`."""
    global _cached
    if _cached is not None:
        return _cached
    try:
        import tiktoken

        _cached = tiktoken.get_encoding("gpt2")
    except Exception as exc:  # ImportError, ConnectionError, ...
        logger.warning("gpt2 tokenizer unavailable, using n_vocab-only stand-in: %s", exc)
        return UnavailableTokenizer(f"{type(exc).__name__}: {exc}")
    return _cached
