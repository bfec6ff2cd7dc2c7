"""Gradient reduction + optimiser step across ranks.

``NcclFallback`` is the transport used across nodes (or whenever peer mapping is impossible):
one NCCL all-reduce over the flat fp32 gradient arena followed by the fused clip + AdamW pass.
The NVLink peer-memory path (``try_create``) replaces both with this repo's own kernels.
"""

from __future__ import annotations

import logging
from typing import Any

import torch
import torch.distributed as dist

from llmtrain.ops import kernels as K

logger = logging.getLogger(__name__)

_live: list[Any] = []


class NcclFallback:
    """torch.distributed (NCCL) all-reduce, then the fused optimiser kernel."""

    name = "nccl"

    def __init__(self, gpt: Any) -> None:
        self._gpt = gpt
        self._world = dist.get_world_size()

    def reduce_and_step(self, m: torch.Tensor, v: torch.Tensor, sumsq: torch.Tensor,
                        norm: torch.Tensor, hyper: dict[str, Any]) -> None:
        g = self._gpt
        dist.all_reduce(g.g32, op=dist.ReduceOp.SUM)
        K.grad_sumsq(g.g32, sumsq)
        K.adamw(g.p32, m, v, g.g32, g.p16, sumsq=sumsq, norm_out=norm,
                grad_scale=1.0 / self._world, **hyper)


def try_create(gpt: Any, m: torch.Tensor, v: torch.Tensor, *, rank: int, world: int,
               device: torch.device, timeout_s: int) -> Any | None:
    """Build the peer-memory communicator, or ``None`` when the ranks cannot map each other."""
    del gpt, m, v, rank, world, device, timeout_s
    return None


def shutdown() -> None:
    """Release peer mappings (called from ``teardown_ddp``)."""
    while _live:
        comm = _live.pop()
        try:
            comm.close()
        except Exception:  # pragma: no cover
            logger.debug("nvl close failed", exc_info=True)
