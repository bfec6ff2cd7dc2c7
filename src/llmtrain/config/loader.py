"""YAML → validated :class:`RunConfig`.

Behavioural parity with `src/llmtrain/config/loader.py:14-65` of the reference: a single
exception type carrying ``message`` / ``details`` / ``errors``, an empty file counts as ``{}``,
and the loader hands back ``(config, path-as-given, absolute-path)``.
"""

from __future__ import annotations

from pathlib import Path
from typing import Any

import yaml
from pydantic import ValidationError

from llmtrain.config.schemas import RunConfig


class ConfigLoadError(Exception):
    """Any failure between "here is a path" and "here is a RunConfig"."""

    def __init__(
        self,
        message: str,
        details: str | None = None,
        errors: list[Any] | None = None,
    ) -> None:
        super().__init__(message)
        self.message = message
        self.details = details
        self.errors = list(errors) if errors else []


def resolve_config_path(config_path: str) -> tuple[str, Path]:
    """Return the path exactly as given plus its absolute, symlink-resolved form."""
    if not (config_path and config_path.strip()):
        raise ConfigLoadError("config path must be a non-empty string")
    expanded = Path(config_path).expanduser()
    if not expanded.is_absolute():
        expanded = Path.cwd() / expanded
    return config_path, expanded.resolve()


def load_yaml_config(config_path: Path) -> Any:
    """Parse a YAML document with the safe loader; ``None`` (empty file) becomes ``{}``."""
    try:
        text = config_path.read_text(encoding="utf-8")
    except OSError as exc:
        raise ConfigLoadError(f"unable to read config file {config_path}: {exc}") from exc
    try:
        document = yaml.safe_load(text)
    except yaml.YAMLError as exc:
        raise ConfigLoadError(f"YAML parse error in {config_path}: {exc}") from exc
    return document if document is not None else {}


def load_and_validate_config(config_path: str) -> tuple[RunConfig, str, Path]:
    """Load, schema-check and freeze a run configuration."""
    given, absolute = resolve_config_path(config_path)
    document = load_yaml_config(absolute)
    if not isinstance(document, dict):
        raise ConfigLoadError(f"top-level config must be a mapping: {absolute}")
    try:
        config = RunConfig.model_validate(document)
    except ValidationError as exc:
        raise ConfigLoadError(
            f"validation failed for {absolute}",
            details=str(exc),
            errors=exc.errors(),
        ) from exc
    return config, given, absolute
