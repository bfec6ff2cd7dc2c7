"""Experiment-tracker contract and the do-nothing implementation.

Signature parity with the reference protocol (`src/llmtrain/tracking/base.py:10-48`).
"""

from __future__ import annotations

from collections.abc import Mapping
from pathlib import Path
from typing import Any, Protocol, runtime_checkable


@runtime_checkable
class Tracker(Protocol):
    """What the trainer and CLI need from a tracking backend."""

    def start_run(self, run_name: str | None = None, *, run_id: str | None = None) -> None: ...

    def log_params(self, params: Mapping[str, Any]) -> None: ...

    def log_metrics(self, metrics: Mapping[str, float], *, step: int | None = None) -> None: ...

    def log_artifact(self, path: str | Path, *, artifact_path: str | None = None) -> None: ...

    def end_run(self) -> None: ...


class NullTracker:
    """Tracking disabled, unavailable, or not rank 0: swallow every call."""

    def start_run(self, run_name: str | None = None, *, run_id: str | None = None) -> None:
        del run_name, run_id

    def log_params(self, params: Mapping[str, Any]) -> None:
        del params

    def log_metrics(self, metrics: Mapping[str, float], *, step: int | None = None) -> None:
        del metrics, step

    def log_artifact(self, path: str | Path, *, artifact_path: str | None = None) -> None:
        del path, artifact_path

    def end_run(self) -> None:
        return None
