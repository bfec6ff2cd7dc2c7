"""Deterministic, filesystem-safe run identifiers.

Format and collision policy follow the reference (`src/llmtrain/utils/run_id.py:29-63`):
``{UTC yyyymmdd_HHMMSS}_{short sha | nogit}_{slug}`` with ``__01``..``__99`` appended when the
directory already exists under ``root_dir``.
"""

from __future__ import annotations

import re
import subprocess
from datetime import datetime, timezone
from pathlib import Path

_NOT_SLUG_CHARS = re.compile(r"[^a-z0-9\-_]+")
_SEPARATOR_RUNS = re.compile(r"[-_]{2,}")
_MAX_SLUG = 40
_MAX_COLLISIONS = 99


def _get_short_git_sha() -> str:
    try:
        proc = subprocess.run(
            ["git", "rev-parse", "--short", "HEAD"],
            check=True,
            capture_output=True,
            text=True,
            timeout=10,
        )
    except (subprocess.CalledProcessError, FileNotFoundError):
        return "nogit"
    return proc.stdout.strip() or "nogit"


def slugify_run_name(name: str) -> str:
    """Lower-case, squash anything outside ``[a-z0-9-_]`` to ``_``, trim, cap at 40 chars."""
    slug = _NOT_SLUG_CHARS.sub("_", name.strip().lower()).strip("-_")
    slug = _SEPARATOR_RUNS.sub("_", slug)
    return slug[:_MAX_SLUG] if slug else "run"


def _first_free(run_id: str, root: Path) -> str:
    if not (root / run_id).exists():
        return run_id
    for n in range(1, _MAX_COLLISIONS + 1):
        candidate = f"{run_id}__{n:02d}"
        if not (root / candidate).exists():
            return candidate
    raise RuntimeError("Run ID collision limit reached (tried __01 through __99).")


def generate_run_id(run_name: str, root_dir: str | Path | None = None) -> str:
    """Build a run id; with ``root_dir`` given, dodge existing directories."""
    stamp = datetime.now(timezone.utc).strftime("%Y%m%d_%H%M%S")
    run_id = "_".join((stamp, _get_short_git_sha(), slugify_run_name(run_name)))
    return run_id if root_dir is None else _first_free(run_id, Path(root_dir))
