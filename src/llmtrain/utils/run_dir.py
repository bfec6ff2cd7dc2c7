"""Run-directory creation and the ``config.yaml`` snapshot.

Same layout and failure semantics as `src/llmtrain/utils/run_dir.py:14-45` in the reference:
``<root>/<run_id>/logs``; an already-existing run directory is an error.
"""

from __future__ import annotations

import shutil
from pathlib import Path

import yaml

from llmtrain.config.schemas import RunConfig
from llmtrain.utils.metadata import atomic_write_text


def create_run_directory(root_dir: str | Path, run_id: str) -> Path:
    """Make ``root/run_id`` (must be new) with a ``logs`` child; roll back on partial failure."""
    root = Path(root_dir)
    root.mkdir(parents=True, exist_ok=True)
    run_path = root / run_id
    run_path.mkdir()  # FileExistsError on duplicates is the contract
    try:
        (run_path / "logs").mkdir()
    except Exception:
        shutil.rmtree(run_path, ignore_errors=True)
        raise
    return run_path


def write_resolved_config(run_dir: str | Path, config: RunConfig) -> Path:
    """Dump the fully-defaulted config in schema order to ``config.yaml``."""
    text = yaml.safe_dump(config.model_dump(), sort_keys=False)
    return atomic_write_text(Path(run_dir) / "config.yaml", text)
