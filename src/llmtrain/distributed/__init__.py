"""Process-group bootstrap: resolve rank info, initialise, tear down.

Contract parity with the reference (`src/llmtrain/distributed/__init__.py:19-153`):
``DDPState`` invariant ``is_main == (rank == 0)``; environment variables win over config
fields; launch by torchrun/K8s is detected through ``RANK``; a second ``setup_ddp`` call
returns the live group's state; ``teardown_ddp`` is a no-op when nothing is initialised.

B200 additions: for ``run.device == "cuda"`` the process binds to ``cuda:LOCAL_RANK`` *before*
the group is created, and the torch process group is only plumbing — backend ``"nccl"`` maps to
``cpu:gloo,cuda:nccl`` (host-side float64 metric tensors keep working) and ``"nvl"`` uses the
same group purely to exchange CUDA-IPC handles for :mod:`llmtrain.parallel.nvl`, whose
peer-memory kernels then carry the gradient traffic.
"""

from __future__ import annotations

import logging
import os
from dataclasses import dataclass
from datetime import timedelta

import torch.distributed as dist

from llmtrain.config.schemas import RunConfig

__all__ = ["DDPState", "setup_ddp", "teardown_ddp"]

logger = logging.getLogger(__name__)

_TORCH_BACKEND = {
    "gloo": "gloo",
    "nccl": "cpu:gloo,cuda:nccl",
    "nvl": "cpu:gloo,cuda:nccl",
}


@dataclass(frozen=True)
class DDPState:
    """Who this process is inside the job."""

    rank: int
    world_size: int
    local_rank: int
    is_main: bool

    def __post_init__(self) -> None:
        if self.is_main is not (self.rank == 0):
            raise ValueError("is_main must be True when rank == 0 and False otherwise")


def _env_int(name: str) -> int | None:
    """Integer value of an environment variable; ``None`` if unset; error if malformed."""
    raw = os.environ.get(name)
    if raw is None:
        return None
    try:
        return int(raw)
    except ValueError:
        raise RuntimeError(f"Env var {name} must be an integer, got: {raw!r}") from None


def _resolve_int(env_name: str, cfg_fallback: int | None, cfg_label: str) -> int:
    """Environment first, then config, else a descriptive ``RuntimeError``."""
    from_env = _env_int(env_name)
    if from_env is not None:
        return from_env
    if cfg_fallback is not None:
        return cfg_fallback
    field = cfg_label.rsplit(".", 1)[-1]
    raise RuntimeError(f"DDP {field} not found in env ({env_name}) or config ({cfg_label})")


def _existing_state(cfg: RunConfig) -> DDPState:
    rank = dist.get_rank()
    local = _env_int("LOCAL_RANK")
    if local is None:
        local = cfg.ddp.local_rank if cfg.ddp.local_rank is not None else rank
    return DDPState(
        rank=rank, world_size=dist.get_world_size(), local_rank=local, is_main=rank == 0
    )


def _bind_cuda_device(local_rank: int) -> None:
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError("run.device is 'cuda' but no CUDA device is visible to this process")
    torch.cuda.set_device(local_rank % torch.cuda.device_count())


def setup_ddp(cfg: RunConfig) -> DDPState:
    """Join (or report) the job-wide process group described by env + ``cfg.ddp``."""
    if dist.is_initialized():
        state = _existing_state(cfg)
        logger.warning(
            "DDP process group already initialised — returning existing state: %s", state
        )
        return state

    under_launcher = "RANK" in os.environ
    rank = _resolve_int("RANK", cfg.ddp.rank, "ddp.rank")
    world_size = _resolve_int("WORLD_SIZE", cfg.ddp.world_size, "ddp.world_size")
    local_rank = _resolve_int("LOCAL_RANK", cfg.ddp.local_rank, "ddp.local_rank")

    init_method = "env://"
    if not under_launcher:
        init_method = cfg.ddp.init_method
        # env:// reads the rendezvous address from the environment: seed it from config.
        for var, value in (
            ("MASTER_ADDR", cfg.ddp.master_addr),
            ("MASTER_PORT", cfg.ddp.master_port),
        ):
            if os.environ.get(var) is None and value is not None:
                os.environ[var] = str(value)

    backend = cfg.ddp.backend
    torch_backend = _TORCH_BACKEND[backend] if cfg.run.device == "cuda" else "gloo"
    if cfg.run.device == "cuda":
        _bind_cuda_device(local_rank)

    logger.info(
        "Initialising DDP process group: rank=%d, world_size=%d, local_rank=%d, "
        "backend=%s, init_method=%s",
        rank,
        world_size,
        local_rank,
        backend,
        init_method,
    )
    dist.init_process_group(
        backend=torch_backend,
        init_method=init_method,
        rank=rank,
        world_size=world_size,
        timeout=timedelta(seconds=cfg.ddp.timeout_sec),
    )
    state = DDPState(rank=rank, world_size=world_size, local_rank=local_rank, is_main=rank == 0)
    logger.info("DDP process group initialised: %s", state)
    return state


def teardown_ddp() -> None:
    """Leave the process group (and release peer-memory mappings) if one is active."""
    try:
        from llmtrain.parallel import nvl

        nvl.shutdown()
    except Exception:  # pragma: no cover - teardown must never mask the real error
        logger.debug("nvl shutdown skipped", exc_info=True)
    if dist.is_initialized():
        logger.info("Destroying DDP process group")
        dist.destroy_process_group()
