"""``meta.json`` — provenance snapshot for a run directory.

Key set mirrors the reference (`src/llmtrain/utils/metadata.py:36-81`); written atomically.
"""

from __future__ import annotations

import json
import os
import platform
import subprocess
import sys
from datetime import datetime, timezone
from pathlib import Path
from typing import Any

DDP_ENV_KEYS = ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")


def _get_git_sha() -> str | None:
    try:
        proc = subprocess.run(
            ["git", "rev-parse", "HEAD"],
            check=True,
            capture_output=True,
            text=True,
            timeout=10,
        )
    except (subprocess.CalledProcessError, FileNotFoundError):
        return "nogit"
    return proc.stdout.strip() or None


def ddp_env_snapshot() -> dict[str, str | None]:
    """The five launcher variables; unset and empty both map to ``None``."""
    return {key: (os.environ.get(key) or None) for key in DDP_ENV_KEYS}


def generate_meta(
    *,
    run_id: str,
    run_name: str,
    config_path: str | None,
    resolved_config_path: str | None,
) -> dict[str, Any]:
    """Collect who/what/where/when for this process."""
    now = datetime.now(timezone.utc).replace(microsecond=0)
    return {
        "meta_version": 1,
        "run_id": run_id,
        "run_name": run_name,
        "created_at": now.strftime("%Y-%m-%dT%H:%M:%SZ"),
        "git_sha": _get_git_sha(),
        "python_version": platform.python_version(),
        "platform": platform.platform(),
        "argv": list(sys.argv),
        "cwd": os.getcwd(),
        "config_path": config_path,
        "resolved_config_path": resolved_config_path,
        "ddp_env": ddp_env_snapshot(),
        "hostname": platform.node(),
        "pid": os.getpid(),
    }


def atomic_write_text(path: Path, text: str) -> Path:
    """Write via ``<name>.tmp`` + rename so readers never see a torn file."""
    tmp = path.with_name(path.name + ".tmp")
    tmp.write_text(text, encoding="utf-8")
    tmp.replace(path)
    return path


def write_meta_json(run_dir: str | Path, meta: dict[str, Any]) -> Path:
    """Persist ``meta`` as ``<run_dir>/meta.json``."""
    body = json.dumps(meta, ensure_ascii=False, indent=2) + "\n"
    return atomic_write_text(Path(run_dir) / "meta.json", body)
