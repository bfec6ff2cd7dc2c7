"""``step_NNNNNN.pt`` checkpoints: write, prune, find, read.

On-disk payload is byte-for-byte the reference's layout (`src/llmtrain/training/checkpoint.py:18-118`,
measured in SURVEY §3.4): ``step``, ``model_state_dict``, ``optimizer_state_dict``,
``scheduler_state_dict``, ``rng_states`` (python / numpy / torch [/ cuda]) and the dumped
``config``.  Two deliberate improvements that do not change the format: the file is written to
``<name>.tmp`` and renamed (a crash mid-save never leaves a truncated "latest" checkpoint —
SURVEY §5.3), and device tensors are staged to host memory before pickling so a checkpoint
written on a B200 loads on a CPU-only box and in the reference.
"""

from __future__ import annotations

import logging
import os
import random
import re
from pathlib import Path
from typing import Any, TypedDict

import numpy as np
import torch

from llmtrain.config.schemas import RunConfig

logger = logging.getLogger(__name__)

_NAME = re.compile(r"^step_(\d+)\.pt$")


class CheckpointPayload(TypedDict):
    """Keys every checkpoint file must contain."""

    step: int
    model_state_dict: dict[str, Any]
    optimizer_state_dict: dict[str, Any]
    scheduler_state_dict: dict[str, Any]
    rng_states: dict[str, Any]
    config: dict[str, Any]


def capture_rng_states() -> dict[str, Any]:
    """Snapshot every generator a resumed run must continue from."""
    states: dict[str, Any] = {
        "python": random.getstate(),
        "numpy": np.random.get_state(),
        "torch": torch.random.get_rng_state(),
    }
    if torch.cuda.is_available():
        states["cuda"] = torch.cuda.get_rng_state_all()
    return states


def restore_rng_states(states: dict[str, Any]) -> None:
    """Inverse of :func:`capture_rng_states` (CUDA part only if CUDA is present here)."""
    random.setstate(states["python"])
    np.random.set_state(states["numpy"])
    torch.random.set_rng_state(states["torch"])
    if "cuda" in states and torch.cuda.is_available():
        cuda_states = states["cuda"]
        if len(cuda_states) == torch.cuda.device_count():
            torch.cuda.set_rng_state_all(cuda_states)


def _to_host(obj: Any, memo: dict[int, torch.Tensor]) -> Any:
    """Deep-copy device tensors to CPU, preserving aliasing (tied weights share storage)."""
    if isinstance(obj, torch.Tensor):
        if obj.device.type == "cpu":
            return obj
        key = id(obj) if obj.numel() == 0 else obj.data_ptr()
        alias = memo.get(key)
        if alias is None or alias.shape != obj.shape:
            alias = obj.detach().to("cpu")
            memo[key] = alias
        return alias
    if isinstance(obj, dict):
        return type(obj)((k, _to_host(v, memo)) for k, v in obj.items())
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_host(v, memo) for v in obj)
    return obj


class CheckpointManager:
    """Owns one ``checkpoints/`` directory."""

    def __init__(self, checkpoint_dir: Path, keep_last_k: int = 3) -> None:
        self._dir = Path(checkpoint_dir)
        self._keep = keep_last_k
        self._dir.mkdir(parents=True, exist_ok=True)

    def _by_step(self) -> list[tuple[int, Path]]:
        found = []
        for path in self._dir.glob("step_*.pt"):
            match = _NAME.match(path.name)
            if match:
                found.append((int(match.group(1)), path))
        return sorted(found)

    def save(
        self,
        step: int,
        model: torch.nn.Module,
        optimizer: torch.optim.Optimizer,
        scheduler: torch.optim.lr_scheduler.LRScheduler,
        config: RunConfig,
    ) -> Path:
        """Write ``step_{step:06d}.pt`` atomically, then drop all but the newest k."""
        memo: dict[int, torch.Tensor] = {}
        payload: CheckpointPayload = {
            "step": step,
            "model_state_dict": _to_host(model.state_dict(), memo),
            "optimizer_state_dict": _to_host(optimizer.state_dict(), memo),
            "scheduler_state_dict": scheduler.state_dict(),
            "rng_states": capture_rng_states(),
            "config": config.model_dump(),
        }
        final = self._dir / f"step_{step:06d}.pt"
        staging = final.with_name(final.name + ".tmp")
        torch.save(payload, staging)
        os.replace(staging, final)
        logger.info("checkpoint: saved step %d to %s", step, final)
        self._prune_old()
        return final

    def load(self, path: Path) -> CheckpointPayload:
        """Read a checkpoint onto the host and verify its key set."""
        logger.info("checkpoint: loading %s", path)
        data = torch.load(path, map_location="cpu", weights_only=False)
        missing = set(CheckpointPayload.__annotations__) - set(data.keys())
        if missing:
            raise ValueError(f"Checkpoint at {path} is missing keys: {missing}")
        return data  # type: ignore[no-any-return]

    def latest_checkpoint(self) -> Path | None:
        """Highest-numbered ``step_*.pt`` here, if any."""
        ordered = self._by_step()
        return ordered[-1][1] if ordered else None

    def _prune_old(self) -> None:
        ordered = self._by_step()
        for _, stale in ordered[: max(0, len(ordered) - self._keep)]:
            stale.unlink()
            logger.debug("checkpoint: pruned %s", stale)
