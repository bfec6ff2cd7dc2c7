"""MLflow adapter (optional dependency).

Call contract as in the reference (`src/llmtrain/tracking/mlflow.py:11-87`): lazy import with
a ``RuntimeError`` when mlflow is absent (the CLI then falls back to ``NullTracker``), params
flattened to dotted keys, metrics coerced to ``float``.
"""

from __future__ import annotations

import json
from collections.abc import Mapping
from pathlib import Path
from typing import Any

_SCALARS = (str, int, float, bool)


def _flatten_params(params: Mapping[str, Any], *, prefix: str = "") -> dict[str, Any]:
    """``{"a": {"b": 1}}`` → ``{"a.b": 1}``; sequences → JSON text; ``None`` → ``"None"``."""
    flat: dict[str, Any] = {}
    stack: list[tuple[str, Mapping[str, Any]]] = [(prefix, params)]
    # Depth-first with an explicit stack, but preserving first-seen key order per level.
    while stack:
        base, mapping = stack.pop()
        nested: list[tuple[str, Mapping[str, Any]]] = []
        for key, value in mapping.items():
            dotted = f"{base}.{key}" if base else str(key)
            if isinstance(value, Mapping):
                nested.append((dotted, value))
            elif isinstance(value, (list, tuple, set)):
                flat[dotted] = json.dumps(list(value), default=str)
            elif value is None:
                flat[dotted] = "None"
            elif isinstance(value, _SCALARS):
                flat[dotted] = value
            else:
                flat[dotted] = str(value)
        stack.extend(reversed(nested))
    return flat


class MLflowTracker:
    """Thin wrapper over the module-level ``mlflow`` fluent API."""

    def __init__(self, *, tracking_uri: str, experiment: str, run_name: str | None = None) -> None:
        try:
            import mlflow  # type: ignore[import-not-found]
        except ModuleNotFoundError as exc:
            raise RuntimeError(
                "MLflowTracker requires the optional 'mlflow' dependency. "
                "Install it with `pip install -e '.[mlflow]'`."
            ) from exc
        self._mlflow = mlflow
        self._tracking_uri = tracking_uri
        self._experiment = experiment
        self._run_name = run_name

    @property
    def active_run_id(self) -> str | None:
        """Id of the run currently open in this process, if any."""
        active = self._mlflow.active_run()
        return None if active is None else active.info.run_id

    def start_run(self, run_name: str | None = None, *, run_id: str | None = None) -> None:
        self._mlflow.set_tracking_uri(self._tracking_uri)
        self._mlflow.set_experiment(self._experiment)
        if run_id is None:
            self._mlflow.start_run(run_name=run_name or self._run_name)
        else:  # attach to a run somebody else created
            self._mlflow.start_run(run_id=run_id)

    def log_params(self, params: Mapping[str, Any]) -> None:
        flat = _flatten_params(params)
        if flat:
            self._mlflow.log_params(flat)

    def log_metrics(self, metrics: Mapping[str, float], *, step: int | None = None) -> None:
        if not metrics:
            return
        values = {key: float(val) for key, val in metrics.items()}
        if step is None:
            self._mlflow.log_metrics(values)
        else:
            self._mlflow.log_metrics(values, step=step)

    def log_artifact(self, path: str | Path, *, artifact_path: str | None = None) -> None:
        self._mlflow.log_artifact(str(path), artifact_path=artifact_path)

    def end_run(self) -> None:
        self._mlflow.end_run()
