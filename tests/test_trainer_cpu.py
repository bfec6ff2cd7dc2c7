"""Trainer on CPU: loop semantics, cadences, checkpoints, resume, tracker contract."""

from __future__ import annotations

import logging
import math
from pathlib import Path
from unittest.mock import Mock

import pytest
import torch

from llmtrain.training import CheckpointManager, Trainer, TrainResult
from llmtrain.training.schedule import warmup_cosine_multiplier
from tests.helpers import TINY_GPT, config


def test_fit_returns_result_and_cpu_reports_zero_peak_memory() -> None:
    result = Trainer(config()).fit()
    assert isinstance(result, TrainResult)
    assert result.final_step == 4 and math.isfinite(result.final_loss)
    assert result.peak_memory == 0.0 and result.total_time >= 0
    assert result.parameter_count == result.trainable_parameter_count > 0
    assert result.first_step_loss is not None and result.resumed_from_step is None


def test_gpt_on_cpu_learns_the_copy_task() -> None:
    cfg = config(model=TINY_GPT,
                 trainer={"max_steps": 80, "micro_batch_size": 4, "lr": 3e-3, "weight_decay": 0.0,
                          "log_every_steps": 40, "eval_every_steps": 80, "save_every_steps": 80})
    result = Trainer(cfg).fit()
    assert result.final_loss < 0.5 * result.first_step_loss
    assert result.final_val_loss is not None and result.final_val_loss < result.first_step_loss


def test_lr_schedule_points() -> None:
    f = warmup_cosine_multiplier(10, 110)
    assert f(0) == 0.0 and f(5) == 0.5 and f(10) == 1.0
    assert f(60) == pytest.approx(0.5) and f(110) == 0.0 and f(500) == 0.0
    assert warmup_cosine_multiplier(0, 10)(0) == 1.0
    assert warmup_cosine_multiplier(10, 10)(9) == 0.9

    trainer = Trainer(config(trainer={"max_steps": 10, "warmup_steps": 4, "lr": 1.0}))
    assert trainer.scheduler.get_last_lr()[0] == 0.0
    assert trainer.scheduler.state_dict()["lr_lambdas"] == [None]  # checkpoint-format parity
    trainer.fit()
    assert trainer.scheduler.get_last_lr()[0] == 0.0


def test_log_eval_and_save_cadences(tmp_path: Path, caplog: pytest.LogCaptureFixture) -> None:
    cfg = config(tmp_path, trainer={"max_steps": 5, "log_every_steps": 2, "eval_every_steps": 2,
                                    "save_every_steps": 2, "extra": {"keep_last_k": 2}})
    tracker = Mock()
    with caplog.at_level(logging.INFO, logger="llmtrain.training.trainer"):
        Trainer(cfg, run_dir=tmp_path / "run", tracker=tracker).fit()
    lines = [r.getMessage() for r in caplog.records]
    assert [m.split()[0] for m in lines if m.startswith("step=")] == ["step=2/5", "step=4/5", "step=5/5"]
    assert [m.split()[0] for m in lines if m.startswith("val_step=")] == [
        "val_step=2/5", "val_step=4/5", "val_step=5/5"]
    assert "tokens_per_sec=" in lines[0] or any("tokens_per_sec=" in m for m in lines)

    tracker.log_params.assert_called_once_with(cfg.model_dump())
    train_calls = [c for c in tracker.log_metrics.call_args_list if "train/loss" in c.args[0]]
    assert [c.kwargs["step"] for c in train_calls] == [2, 4, 5]
    assert set(train_calls[0].args[0]) == {"train/loss", "train/lr", "train/tokens_per_sec",
                                           "train/step_time_sec", "train/tokens_total"}
    val_calls = [c for c in tracker.log_metrics.call_args_list if "val/loss" in c.args[0]]
    assert [c.kwargs["step"] for c in val_calls] == [2, 4, 5]

    names = sorted(p.name for p in (tmp_path / "run" / "checkpoints").glob("*.pt"))
    assert names == ["step_000004.pt", "step_000005.pt"]  # saved at 2,4,5; keep_last_k=2
    assert not list((tmp_path / "run" / "checkpoints").glob("*.tmp"))


def test_deferred_log_lines_keep_their_order_values_and_steps(tmp_path: Path) -> None:
    """A train log line is written one step late (after the next step has been enqueued) but is
    still the line of ITS step: before that step's validation line, with that step's learning
    rate and running token total, and every step is logged exactly once."""
    cfg = config(tmp_path, trainer={"max_steps": 6, "log_every_steps": 1, "eval_every_steps": 3,
                                    "save_every_steps": 100, "warmup_steps": 2, "lr": 1e-2})
    tracker = Mock()
    trainer = Trainer(cfg, run_dir=tmp_path / "run", tracker=tracker)
    trainer.fit()
    order = [("val" if "val/loss" in c.args[0] else "train", c.kwargs["step"])
             for c in tracker.log_metrics.call_args_list
             if "val/loss" in c.args[0] or "train/loss" in c.args[0]]
    assert order == [("train", 1), ("train", 2), ("train", 3), ("val", 3), ("train", 4),
                     ("train", 5), ("train", 6), ("val", 6)]
    train = {c.kwargs["step"]: c.args[0] for c in tracker.log_metrics.call_args_list
             if "train/loss" in c.args[0]}
    per_step = train[1]["train/tokens_total"]
    assert per_step > 0 and [train[k]["train/tokens_total"] for k in range(1, 7)] == [
        k * per_step for k in range(1, 7)]
    # warm-up over 2 steps: the line of step 1 carries the rate in force after step 1, and so on
    f = warmup_cosine_multiplier(2, 6)
    assert [train[k]["train/lr"] for k in range(1, 7)] == pytest.approx(
        [1e-2 * f(k) for k in range(1, 7)])


def test_checkpoint_payload_and_manager(tmp_path: Path) -> None:
    cfg = config(tmp_path, model=TINY_GPT, trainer={"max_steps": 2, "save_every_steps": 1})
    trainer = Trainer(cfg, run_dir=tmp_path / "r")
    trainer.fit()
    mgr = CheckpointManager(tmp_path / "r" / "checkpoints")
    latest = mgr.latest_checkpoint()
    assert latest is not None and latest.name == "step_000002.pt"
    payload = mgr.load(latest)
    assert set(payload) == {"step", "model_state_dict", "optimizer_state_dict",
                            "scheduler_state_dict", "rng_states", "config"}
    assert payload["step"] == 2 and payload["config"] == cfg.model_dump()
    assert {"python", "numpy", "torch"} <= set(payload["rng_states"])
    sd = payload["model_state_dict"]
    assert sd["lm_head.weight"].data_ptr() == sd["token_embedding.weight"].data_ptr()
    assert "blocks.0.attn.causal_mask" in sd
    group = payload["optimizer_state_dict"]["param_groups"][0]
    assert group["lr"] >= 0 and group["betas"] == (0.9, 0.999) and group["weight_decay"] == 0.1
    assert len(group["params"]) == 2 + 12 * 2 + 2
    torch.save({"step": 1}, tmp_path / "broken.pt")
    with pytest.raises(ValueError, match="missing keys"):
        mgr.load(tmp_path / "broken.pt")


def test_resume_equals_uninterrupted_run(tmp_path: Path) -> None:
    torch.use_deterministic_algorithms(True)
    try:
        base = dict(model=TINY_GPT, trainer={"max_steps": 8, "micro_batch_size": 2,
                                              "grad_accum_steps": 2, "save_every_steps": 4,
                                              "lr": 1e-3, "log_every_steps": 100,
                                              "eval_every_steps": 100})
        straight = Trainer(config(tmp_path, **base), run_dir=tmp_path / "a")
        want = straight.fit()
        # a fresh process-equivalent: new trainer, different init, everything comes from step 4
        torch.manual_seed(12345)
        second = Trainer(config(tmp_path, **base), run_dir=tmp_path / "c")
        got = second.fit(resume_from=tmp_path / "a" / "checkpoints" / "step_000004.pt")
        assert got.resumed_from_step == 4 and got.first_step_loss is None
        assert abs(got.final_loss - want.final_loss) < 1e-5
    finally:
        torch.use_deterministic_algorithms(False)


def test_resume_path_resolution_and_config_mismatch_warning(
    tmp_path: Path, caplog: pytest.LogCaptureFixture
) -> None:
    cfg = config(tmp_path / "runs", trainer={"max_steps": 2, "save_every_steps": 2})
    Trainer(cfg, run_dir=tmp_path / "runs" / "rid").fit()
    trainer = Trainer(config(tmp_path / "runs", trainer={"max_steps": 3, "save_every_steps": 2}))
    assert trainer._resolve_resume_path("rid").name == "step_000002.pt"  # by run id
    ckpt = tmp_path / "runs" / "rid" / "checkpoints" / "step_000002.pt"
    assert trainer._resolve_resume_path(ckpt) == ckpt
    assert trainer._resolve_resume_path(ckpt.parent) == ckpt
    with pytest.raises(FileNotFoundError, match="does not exist"):
        trainer._resolve_resume_path(tmp_path / "nope.pt")
    with pytest.raises(FileNotFoundError, match="Checkpoint directory"):
        trainer._resolve_resume_path("unknown-run")
    with caplog.at_level(logging.WARNING, logger="llmtrain.training.trainer"):
        result = trainer.fit(resume_from="rid")
    assert result.resumed_from_step == 2 and result.final_step == 3
    assert any("config mismatch" in r.getMessage() for r in caplog.records)


def test_dry_run_steps_and_early_stop() -> None:
    from llmtrain.training.dry_run import DEFAULT_DRY_RUN_STEPS, run_dry_run

    out = run_dry_run(config(trainer={"max_steps": 50, "micro_batch_size": 1}))
    assert out.steps_executed == DEFAULT_DRY_RUN_STEPS == 5
    assert (out.resolved_model_adapter, out.resolved_data_module) == ("dummy_gpt", "dummy_text")
    assert run_dry_run(config(trainer={"max_steps": 2})).steps_executed == 2
    short = run_dry_run(config(trainer={"max_steps": 1, "micro_batch_size": 1}))
    assert short.steps_executed == 1


def test_engine_selection_rules() -> None:
    from llmtrain.engine.torch_engine import TorchEngine

    assert isinstance(Trainer(config())._engine, TorchEngine)
    with pytest.raises(ValueError, match="engine must be one of"):
        Trainer(config(trainer={"extra": {"engine": "tpu"}}))
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CUDA device"):
            Trainer(config(run={"device": "cuda"}))


def test_step_scalars_wait_once_for_their_event_and_keep_their_value() -> None:
    """What the B200 engine's ``train_step`` returns: reading waits for the step's own event (not
    for the stream), once, and the value survives the ring slot being rewritten."""
    from llmtrain.engine.b200 import StepScalars

    class Event:
        waits = 0

        def synchronize(self) -> None:
            Event.waits += 1

    row = torch.tensor([1.5, 2.5])
    scalars = StepScalars(row, Event())  # type: ignore[arg-type]
    assert Event.waits == 0  # nothing is read until somebody asks
    assert float(scalars) == 1.5 and scalars.grad_norm == 2.5 and scalars.item() == 1.5
    assert Event.waits == 1
    row[0] = 9.0  # the slot is reused by a later step
    assert float(scalars) == 1.5
    assert not torch.is_tensor(scalars)  # the trainer's interval sums it as a host number

