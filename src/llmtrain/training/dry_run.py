"""``--dry-run``: build everything, push a handful of batches through the loss, touch nothing.

Same contract as the reference (`src/llmtrain/training/dry_run.py:15-87`): always on the CPU,
model in train mode under ``no_grad``, ``min(5, max_steps)`` steps, stops early when the loader
runs dry, one ``Dry-run step i/n loss=… step_time_ms=…`` log line per step.
"""

from __future__ import annotations

import logging
import time
from dataclasses import dataclass

import torch

from llmtrain.config.schemas import RunConfig
from llmtrain.engine.torch_engine import move_batch
from llmtrain.registry import initialize_registries
from llmtrain.registry.data import get_data_module
from llmtrain.registry.models import get_model_adapter

DEFAULT_DRY_RUN_STEPS = 5


@dataclass(frozen=True)
class DryRunResult:
    resolved_model_adapter: str
    resolved_data_module: str
    steps_executed: int


def run_dry_run(cfg: RunConfig, *, logger: logging.Logger | None = None) -> DryRunResult:
    """Forward-only sanity pass over at most ``DEFAULT_DRY_RUN_STEPS`` training batches."""
    log = logger if logger is not None else logging.getLogger(__name__)
    initialize_registries()
    adapter = get_model_adapter(cfg.model.name)()
    data_module = get_data_module(cfg.data.name)()

    cpu = torch.device("cpu")
    model = adapter.build_model(cfg).to(cpu).train()
    data_module.setup(cfg, tokenizer=adapter.build_tokenizer(cfg))

    planned = min(max(DEFAULT_DRY_RUN_STEPS, 1), cfg.trainer.max_steps)
    done = 0
    with torch.no_grad():
        for done_so_far, batch in enumerate(data_module.train_dataloader(), start=1):
            if done_so_far > planned:
                break
            tick = time.perf_counter()
            loss, metrics = adapter.compute_loss(model, move_batch(batch, cpu))
            elapsed_ms = 1e3 * (time.perf_counter() - tick)
            done = done_so_far
            log.info(
                "Dry-run step %s/%s loss=%.4f step_time_ms=%.2f",
                done,
                planned,
                metrics.get("loss", float(loss.item())),
                elapsed_ms,
            )
    return DryRunResult(
        resolved_model_adapter=cfg.model.name,
        resolved_data_module=cfg.data.name,
        steps_executed=done,
    )
