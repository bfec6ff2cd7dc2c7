"""Generic name → class registry used by both plugin kinds."""

from __future__ import annotations

from collections.abc import Callable
from typing import Generic, TypeVar

C = TypeVar("C")


class Registry(Generic[C]):
    """Process-global, append-only plugin table.

    ``kind`` is the human label used in messages (``"Model adapter"`` / ``"Data module"``) and
    ``error_type`` is the exception class raised, so that each plugin family keeps its own
    distinct ``RegistryError`` (the CLI catches both; reference `src/llmtrain/cli.py:18-21,242`).
    """

    def __init__(self, kind: str, error_type: type[ValueError]) -> None:
        self._kind = kind
        self._error = error_type
        self._table: dict[str, type] = {}

    def _clean(self, name: str) -> str:
        cleaned = name.strip()
        if not cleaned:
            raise self._error("Registry name must be non-empty.")
        return cleaned

    def _listing(self) -> str:
        return ", ".join(sorted(self._table)) or "none"

    def register(self, name: str) -> Callable[[type[C]], type[C]]:
        key = self._clean(name)

        def _decorate(cls: type[C]) -> type[C]:
            if key in self._table:
                raise self._error(
                    f"{self._kind} '{key}' is already registered. Available: {self._listing()}."
                )
            self._table[key] = cls
            return cls

        return _decorate

    def get(self, name: str) -> type:
        key = self._clean(name)
        try:
            return self._table[key]
        except KeyError:
            raise self._error(
                f"Unknown {self._kind.lower()} '{key}'. Available: {self._listing()}."
            ) from None

    def names(self) -> list[str]:
        return sorted(self._table)
