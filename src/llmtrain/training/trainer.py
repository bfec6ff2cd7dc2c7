"""Step-based training loop with gradient accumulation, cadenced logging / eval / checkpoints.

Observable behaviour follows the reference trainer (`src/llmtrain/training/trainer.py:56-548`):
constructor wiring, metric key names (``train/*``, ``train/*_rank_{r}``, ``val/*``), log-line
formats, save / log / eval cadences (each also fires on the final step), resume semantics
(skip ``step * accum`` batches in single-process mode, ``first_step_loss`` stays ``None``),
rank-0-only side effects, and the ``TrainResult`` fields.

What differs is *who executes a step*: the loop delegates to a
:class:`~llmtrain.engine.base.TrainEngine`.  On CPU that is torch autograd + DDP + AdamW (the
reference's own sequence); on a B200 it is the native runtime, where the loss stays on the
device and is only read back when a cadence needs it and ``peak_memory`` is
the real high-water mark instead of the reference's hard-coded ``0.0``.
"""

from __future__ import annotations

import logging
import os
import sys
import time
from pathlib import Path
from typing import Any

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel

from llmtrain.config.schemas import RunConfig
from llmtrain.distributed import DDPState
from llmtrain.engine.base import Batch, TrainEngine
from llmtrain.engine.torch_engine import TorchEngine, move_batch
from llmtrain.registry import initialize_registries
from llmtrain.registry.data import get_data_module
from llmtrain.registry.models import get_model_adapter
from llmtrain.tracking import NullTracker, Tracker
from llmtrain.training.checkpoint import (
    CheckpointManager,
    CheckpointPayload,
    restore_rng_states,
)
from llmtrain.training.result import TrainResult
from llmtrain.training.schedule import build_scheduler

__all__ = ["TrainResult", "Trainer"]

logger = logging.getLogger(__name__)


def _move_batch(batch: Batch, device: torch.device) -> Batch:
    return move_batch(batch, device)


class _Interval:
    """Running sums between two log lines."""

    def __init__(self) -> None:
        self.reset()

    def reset(self) -> None:
        self.losses: list[Any] = []
        self.tokens = 0
        self.started = time.perf_counter()

    def add(self, loss: Any, tokens: int) -> None:
        self.losses.append(loss)
        self.tokens += tokens

    @property
    def steps(self) -> int:
        return len(self.losses)

    def loss_sum(self) -> float:
        """Host value of the summed step losses (a single device read when they are tensors)."""
        host = sum(float(v) for v in self.losses if not torch.is_tensor(v))
        on_device = [v.detach().float().reshape(()) for v in self.losses if torch.is_tensor(v)]
        if on_device:
            host += float(torch.stack(on_device).sum())
        return host


class Trainer:
    """One process, one device; optional data parallelism across processes."""

    def __init__(
        self,
        cfg: RunConfig,
        *,
        run_dir: Path | None = None,
        tracker: Tracker | None = None,
        ddp_state: DDPState | None = None,
    ) -> None:
        self._cfg = cfg
        self._tracker: Tracker = tracker if tracker is not None else NullTracker()
        self._ddp_state = ddp_state
        initialize_registries()
        self._adapter = get_model_adapter(cfg.model.name)()
        data_module = get_data_module(cfg.data.name)()

        model = self._adapter.build_model(cfg)
        tokenizer = self._adapter.build_tokenizer(cfg)
        data_module.setup(cfg, tokenizer=tokenizer)
        self._train_loader = data_module.train_dataloader()
        self._val_loader = data_module.val_dataloader()

        self._device = self._resolve_device()
        model = model.to(self._device)
        self._engine = self._build_engine(model)
        self._scheduler = build_scheduler(
            self._engine.optimizer, cfg.trainer.warmup_steps, cfg.trainer.max_steps
        )
        self._ckpt_mgr: CheckpointManager | None = None
        if run_dir is not None:
            keep = int(cfg.trainer.extra.get("keep_last_k", 3))
            self._ckpt_mgr = CheckpointManager(Path(run_dir) / "checkpoints", keep_last_k=keep)

    # ------------------------------------------------------------------ construction helpers
    def _resolve_device(self) -> torch.device:
        kind = self._cfg.run.device
        if kind != "cuda":
            return torch.device(kind)
        if not torch.cuda.is_available():
            raise RuntimeError("run.device is 'cuda' but torch sees no CUDA device")
        # without a DDP state: the device the caller made current (0 in a fresh process), so a
        # trainer built inside a multi-GPU program stays on that process's GPU
        local = (self._ddp_state.local_rank if self._ddp_state is not None
                 else torch.cuda.current_device())
        index = local % torch.cuda.device_count()
        torch.cuda.set_device(index)
        return torch.device("cuda", index)

    def _build_engine(self, model: torch.nn.Module) -> TrainEngine:
        cfg = self._cfg
        choice = str(cfg.trainer.extra.get("engine", "auto"))
        if choice not in ("auto", "torch", "b200"):
            raise ValueError("trainer.extra.engine must be one of: auto, torch, b200")
        if self._device.type == "cuda" and choice in ("auto", "b200"):
            from llmtrain.engine.b200 import B200TrainEngine, supports

            blocker = None
            if not supports(model):
                blocker = f"model {type(model).__name__} is outside the kernels' shape coverage"
            if blocker is None:
                return B200TrainEngine(
                    model, cfg, device=self._device, ddp_state=self._ddp_state
                )
            if choice == "b200":
                raise ValueError(f"engine 'b200' unavailable: {blocker}")
            logger.warning("native B200 engine not used (%s); training on torch CUDA ops", blocker)
        return TorchEngine(
            model,
            self._adapter,
            device=self._device,
            lr=cfg.trainer.lr,
            weight_decay=cfg.trainer.weight_decay,
            max_grad_norm=cfg.trainer.max_grad_norm,
            data_parallel=self._is_ddp_active,
            find_unused_parameters=cfg.ddp.find_unused_parameters,
        )

    # ------------------------------------------------------------------ small accessors
    @property
    def _model(self) -> torch.nn.Module:
        return self._engine.model

    @property
    def _optimizer(self) -> torch.optim.Optimizer:
        return self._engine.optimizer

    @property
    def scheduler(self) -> torch.optim.lr_scheduler.LambdaLR:
        """The LR scheduler (exposed so callers can inspect the curve)."""
        return self._scheduler

    @property
    def _is_main(self) -> bool:
        return self._ddp_state is None or self._ddp_state.is_main

    @property
    def _raw_model(self) -> torch.nn.Module:
        model = self._engine.model
        return model.module if isinstance(model, DistributedDataParallel) else model

    @property
    def _rank(self) -> int:
        return 0 if self._ddp_state is None else self._ddp_state.rank

    @property
    def _world_size(self) -> int:
        return 1 if self._ddp_state is None else self._ddp_state.world_size

    @property
    def _is_ddp_active(self) -> bool:
        return self._ddp_state is not None and self._ddp_state.world_size > 1

    # ------------------------------------------------------------------ cross-rank scalars
    # float64 host tensors through ``dist.all_reduce`` / ``dist.all_gather`` — looked up on the
    # module at call time so tests (and alternative transports) can substitute them.
    def _reduce_metrics(self, **scalars: float) -> dict[str, float]:
        """Element-wise SUM over ranks; identity without data parallelism."""
        if not self._is_ddp_active:
            return dict(scalars)
        names = list(scalars)
        packed = torch.tensor([scalars[n] for n in names], dtype=torch.float64)
        dist.all_reduce(packed, op=dist.ReduceOp.SUM)
        return {n: float(v) for n, v in zip(names, packed.tolist(), strict=True)}

    def _gather_scalars(self, **scalars: float) -> list[dict[str, float]] | None:
        """Per-rank values, delivered to rank 0 only (``None`` elsewhere)."""
        if not self._is_ddp_active:
            return [dict(scalars)]
        if not (dist.is_available() and dist.is_initialized()):
            # Data parallelism was requested but there is no group: never let a non-main rank
            # obtain a list (it would start logging metrics).
            return [dict(scalars)] if self._is_main else None
        names = list(scalars)
        mine = torch.tensor([scalars[n] for n in names], dtype=torch.float64)
        slots = [torch.zeros_like(mine) for _ in range(self._world_size)]
        dist.all_gather(slots, mine)
        if not self._is_main:
            return None
        return [{n: float(v) for n, v in zip(names, slot.tolist(), strict=True)} for slot in slots]

    # ------------------------------------------------------------------ checkpoint restore
    def restore(self, payload: CheckpointPayload) -> int:
        """Load model / optimiser / scheduler / RNG state; returns the checkpoint's step."""
        self._raw_model.load_state_dict(payload["model_state_dict"])
        self._optimizer.load_state_dict(payload["optimizer_state_dict"])
        self._scheduler.load_state_dict(payload["scheduler_state_dict"])
        restore_rng_states(payload["rng_states"])
        self._engine.after_restore()
        step = int(payload["step"])
        logger.info("trainer: restored state from step %d", step)
        return step

    def _resolve_resume_path(self, resume_from: str | Path) -> Path:
        """Accept a ``.pt`` file, a checkpoint directory, or a run id under ``root_dir``."""

        def newest_in(folder: Path) -> Path:
            latest = CheckpointManager(folder, keep_last_k=1).latest_checkpoint()
            if latest is None:
                raise FileNotFoundError(f"No checkpoints found in {folder}")
            return latest

        given = Path(resume_from)
        if given.is_file():
            return given
        if given.is_dir():
            return newest_in(given)
        if given.exists():
            raise FileNotFoundError(f"Resume path {given} is not a file or directory")
        if given.suffix == ".pt":
            raise FileNotFoundError(f"Checkpoint file {given} does not exist")
        folder = Path(self._cfg.output.root_dir) / str(resume_from) / "checkpoints"
        if not folder.exists():
            raise FileNotFoundError(f"Checkpoint directory {folder} does not exist")
        return newest_in(folder)

    # ------------------------------------------------------------------ validation
    def _evaluate(self) -> tuple[dict[str, float], dict[str, float]] | None:
        """Token-weighted validation loss: ``(this rank, all ranks)``; ``None`` without data."""
        if self._val_loader is None:
            return None
        model = self._model
        was_training = model.training
        model.eval()
        loss_sum, tokens = 0.0, 0
        with torch.no_grad():
            for batch in self._val_loader:
                n = int(batch["input_ids"].numel())
                loss_sum += self._engine.eval_loss(batch) * n
                tokens += n
        if was_training:
            model.train()
        if tokens == 0:
            return {}, {}
        local = {"val/loss": loss_sum / tokens}
        if not self._is_ddp_active:
            return local, dict(local)
        total = self._reduce_metrics(loss_sum=loss_sum, tok_count=float(tokens))
        pooled = total["loss_sum"] / total["tok_count"] if total["tok_count"] > 0 else 0.0
        return local, {"val/loss": pooled}

    # ------------------------------------------------------------------ logging blocks
    def _log_train_interval(self, step: int, max_steps: int, window: _Interval,
                            tokens_local_total: int, tokens_global_total: int,
                            lr_now: float) -> int:
        # Read the losses back FIRST: on CUDA that is the synchronisation point, and the interval
        # must not end while the interval's kernels are still in flight.
        loss_sum = window.loss_sum()
        elapsed = time.perf_counter() - window.started
        mean_loss = loss_sum / window.steps
        step_time = elapsed / window.steps
        rate = window.tokens / elapsed if elapsed > 0 else 0.0
        extra = self._engine.extra_metrics(rate, window.losses[-1])

        if self._is_ddp_active:
            rows = self._gather_scalars(
                avg_loss=mean_loss,
                lr=lr_now,
                tokens_per_sec=rate,
                step_time=step_time,
                tokens_total=float(tokens_local_total),
            )
            for r, row in enumerate(rows or ()):
                self._tracker.log_metrics(
                    {
                        f"train/loss_rank_{r}": row["avg_loss"],
                        f"train/lr_rank_{r}": row["lr"],
                        f"train/tokens_per_sec_rank_{r}": row["tokens_per_sec"],
                        f"train/step_time_sec_rank_{r}": row["step_time"],
                        f"train/tokens_total_rank_{r}": row["tokens_total"],
                    },
                    step=step,
                )
            pooled = self._reduce_metrics(
                loss_sum=loss_sum, steps=float(window.steps), tokens=float(window.tokens)
            )
            if self._is_main:
                tokens_global_total += int(pooled["tokens"])
                self._tracker.log_metrics(
                    {
                        "train/loss": pooled["loss_sum"] / pooled["steps"],
                        "train/lr": lr_now,
                        "train/tokens_per_sec": pooled["tokens"] / elapsed if elapsed > 0 else 0.0,
                        "train/tokens_total": float(tokens_global_total),
                        "train/step_time_sec": step_time,
                        **extra,
                    },
                    step=step,
                )
        elif self._is_main:
            self._tracker.log_metrics(
                {
                    "train/loss": mean_loss,
                    "train/lr": lr_now,
                    "train/tokens_per_sec": rate,
                    "train/step_time_sec": step_time,
                    "train/tokens_total": float(tokens_local_total),
                    **extra,
                },
                step=step,
            )
        logger.info(
            "step=%d/%d  loss=%.4f  lr=%.6e  tokens_per_sec=%.1f  step_time=%.4fs",
            step,
            max_steps,
            mean_loss,
            lr_now,
            rate,
            step_time,
        )
        window.reset()
        return tokens_global_total

    # ------------------------------------------------------------------ the loop
    def fit(
        self,
        *,
        max_steps_override: int | None = None,
        resume_from: str | Path | None = None,
    ) -> TrainResult:
        """Train until ``max_steps`` optimiser updates have been applied."""
        cfg = self._cfg.trainer
        t_fit = time.perf_counter()
        trace_fit = os.environ.get("LLMTRAIN_FIT_TRACE") is not None  # host-side phase times
        self._model.train()
        max_steps = cfg.max_steps if max_steps_override is None else max_steps_override
        accum = cfg.grad_accum_steps

        loader_iter = iter(self._train_loader)

        def next_batch() -> Batch:
            nonlocal loader_iter
            try:
                return next(loader_iter)
            except StopIteration:
                loader_iter = iter(self._train_loader)
                return next(loader_iter)

        first_step = 1
        resumed_from_step: int | None = None
        if resume_from is not None:
            path = self._resolve_resume_path(resume_from)
            payload = CheckpointManager(path.parent, keep_last_k=1).load(path)
            if payload["config"] != self._cfg.model_dump():
                logger.warning(
                    "checkpoint: config mismatch detected; using current config for resume"
                )
            resumed_from_step = self.restore(payload)
            first_step = resumed_from_step + 1
            if first_step > max_steps:
                logger.info(
                    "trainer: resume step %d >= max_steps %d; no further steps",
                    resumed_from_step,
                    max_steps,
                )

        if self._is_main:
            self._tracker.log_params(self._cfg.model_dump())
        params = list(self._raw_model.parameters())
        parameter_count = sum(p.numel() for p in params)
        trainable_parameter_count = sum(p.numel() for p in params if p.requires_grad)

        t_start = time.perf_counter()
        if resumed_from_step and not self._is_ddp_active:
            # Re-walk the loader to roughly the position the interrupted run had reached.
            # Not done under DDP: every rank holds a different shard (see reference comment,
            # `src/llmtrain/training/trainer.py:336-339`).
            for _ in range(resumed_from_step * accum):
                next_batch()

        first_step_obj: Any = None
        final_val_loss: float | None = None
        final_val_metrics: dict[str, float] | None = None
        last_loss: Any = 0.0
        tokens_local_total = 0
        tokens_global_total = 0
        window = _Interval()
        # A log line is written one step late: the interval's scalars are read back (and, under
        # DDP, exchanged between the ranks) after the NEXT step has been enqueued, so the device
        # never waits for the host's logging.  Values, step numbers and the order relative to
        # checkpoints and evaluations are those of immediate logging.
        pending: tuple[int, _Interval, int, float] | None = None

        def flush_log() -> None:
            nonlocal pending, tokens_global_total
            if pending is None:
                return
            at_step, interval, tokens_then, lr_then = pending
            pending = None
            tokens_global_total = self._log_train_interval(
                at_step, max_steps, interval, tokens_then, tokens_global_total, lr_then
            )
            window.started = time.perf_counter()  # the next interval starts where this one ended

        t_loop = time.perf_counter()
        t_first = t_loop
        trace_rows: list[tuple[float, float, float, Any]] = []
        for step in range(first_step, max_steps + 1):
            t_call = time.perf_counter()
            last_loss, step_tokens = self._engine.train_step(next_batch, accum)
            t_enq = time.perf_counter()
            if step == first_step:
                t_first = t_enq
            flush_log()  # the previous interval: the device is busy with this step meanwhile
            if trace_fit:
                trace_rows.append((t_call - t_loop, t_enq - t_loop, time.perf_counter() - t_loop,
                                   getattr(last_loss, "timing_event", None)))
            self._scheduler.step()
            tokens_local_total += step_tokens
            if step == 1:
                first_step_obj = last_loss  # read at the end: no host sync in the first step
            is_final = step == max_steps

            if (
                self._ckpt_mgr is not None
                and self._is_main
                and (step % cfg.save_every_steps == 0 or is_final)
            ):
                self._engine.before_checkpoint()
                self._ckpt_mgr.save(
                    step, self._raw_model, self._optimizer, self._scheduler, self._cfg
                )

            window.add(last_loss, step_tokens)
            if step % cfg.log_every_steps == 0 or is_final:
                pending = (step, window, tokens_local_total,
                           float(self._scheduler.get_last_lr()[0]))
                window = _Interval()
                if is_final:
                    flush_log()

            if step % cfg.eval_every_steps == 0 or is_final:
                flush_log()
                outcome = self._evaluate()
                if outcome is None:
                    continue  # no validation data configured
                local_val, global_val = outcome
                if self._is_ddp_active:
                    rows = self._gather_scalars(val_loss=local_val.get("val/loss", 0.0))
                    for r, row in enumerate(rows or ()):
                        self._tracker.log_metrics({f"val/loss_rank_{r}": row["val_loss"]}, step=step)
                if self._is_main:
                    self._tracker.log_metrics(global_val, step=step)
                seen = global_val if self._is_main else local_val
                if seen:
                    final_val_metrics = seen
                    if seen.get("val/loss") is not None:
                        final_val_loss = seen["val/loss"]
                    rendered = "  ".join(f"{k}={v:.4f}" for k, v in sorted(seen.items()))
                    logger.info("val_step=%d/%d  %s", step, max_steps, rendered)

        if trace_fit:
            now = time.perf_counter()
            print(f"fit trace: prologue {1e3 * (t_loop - t_fit):.2f} ms, first train_step call "
                  f"{1e3 * (t_first - t_loop):.2f} ms, loop {1e3 * (now - t_loop):.2f} ms",
                  file=sys.stderr, flush=True)
            prev = None
            for i, (a, b, c, ev) in enumerate(trace_rows):
                gap = "" if prev is None or ev is None else f"  device step {prev.elapsed_time(ev):.2f} ms"
                print(f"fit trace: step {i + 1}: train_step called at {1e3 * a:.2f} ms, enqueued by "
                      f"{1e3 * b:.2f}, previous log flushed by {1e3 * c:.2f}{gap}", file=sys.stderr)
                prev = ev if ev is not None else prev
        return TrainResult(
            final_step=max_steps,
            final_loss=float(last_loss),
            final_val_loss=final_val_loss,
            total_time=time.perf_counter() - t_start,
            peak_memory=self._engine.peak_memory_bytes(),
            val_metrics=final_val_metrics,
            first_step_loss=None if first_step_obj is None else float(first_step_obj),
            resumed_from_step=resumed_from_step,
            parameter_count=parameter_count,
            trainable_parameter_count=trainable_parameter_count,
        )
