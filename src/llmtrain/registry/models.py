"""Model-adapter registry (reference API: `src/llmtrain/registry/models.py:13-54`)."""

from __future__ import annotations

from collections.abc import Callable
from typing import TypeVar

from llmtrain.models.base import ModelAdapter
from llmtrain.registry._core import Registry

T = TypeVar("T", bound=ModelAdapter)


class RegistryError(ValueError):
    """Model-registry failure (unknown, duplicate or empty name)."""


_REGISTRY: Registry[ModelAdapter] = Registry("Model adapter", RegistryError)
_MODEL_REGISTRY = _REGISTRY._table  # plain-dict view kept for introspection/tests


def register_model(name: str) -> Callable[[type[T]], type[T]]:
    """Class decorator: make ``cls`` constructible from ``model.name == name``."""
    return _REGISTRY.register(name)  # type: ignore[return-value]


def get_model_adapter(name: str) -> type[ModelAdapter]:
    """Look up an adapter class; the error message lists what is available."""
    return _REGISTRY.get(name)


def available_model_adapters() -> list[str]:
    """Sorted registered names."""
    return _REGISTRY.names()
