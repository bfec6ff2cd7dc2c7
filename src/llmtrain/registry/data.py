"""Data-module registry (reference API: `src/llmtrain/registry/data.py:13-54`)."""

from __future__ import annotations

from collections.abc import Callable
from typing import TypeVar

from llmtrain.data.base import DataModule
from llmtrain.registry._core import Registry

T = TypeVar("T", bound=DataModule)


class RegistryError(ValueError):
    """Data-registry failure (unknown, duplicate or empty name)."""


_REGISTRY: Registry[DataModule] = Registry("Data module", RegistryError)
_DATA_REGISTRY = _REGISTRY._table


def register_data_module(name: str) -> Callable[[type[T]], type[T]]:
    """Class decorator: make ``cls`` constructible from ``data.name == name``."""
    return _REGISTRY.register(name)  # type: ignore[return-value]


def get_data_module(name: str) -> type[DataModule]:
    """Look up a data-module class; the error message lists what is available."""
    return _REGISTRY.get(name)


def available_data_modules() -> list[str]:
    """Sorted registered names."""
    return _REGISTRY.names()
