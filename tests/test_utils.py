"""run ids, run dirs, metadata, logging, summaries."""

from __future__ import annotations

import json
import logging
import subprocess
from datetime import datetime
from pathlib import Path

import pytest
import yaml

from llmtrain.training.result import TrainResult
from llmtrain.utils import run_id as run_id_mod
from llmtrain.utils.logging import JsonFormatter, configure_logging
from llmtrain.utils.metadata import generate_meta, write_meta_json
from llmtrain.utils.run_dir import create_run_directory, write_resolved_config
from llmtrain.utils.summary import format_run_summary
from tests.helpers import config


class _Frozen(datetime):
    @classmethod
    def now(cls, tz=None):  # noqa: ANN001, ANN206
        return datetime(2031, 5, 6, 7, 8, 9, tzinfo=tz)


@pytest.mark.parametrize(
    "raw,slug",
    [("My Run!", "my_run"), ("  --a__b--  ", "a_b"), ("", "run"), ("Ünï", "n"), ("x" * 60, "x" * 40),
     ("keep-dash_and_under", "keep-dash_and_under")],
)
def test_slugify(raw: str, slug: str) -> None:
    assert run_id_mod.slugify_run_name(raw) == slug


def test_run_id_format_and_collisions(monkeypatch: pytest.MonkeyPatch, tmp_path: Path) -> None:
    monkeypatch.setattr(run_id_mod, "datetime", _Frozen)
    monkeypatch.setattr(run_id_mod, "_get_short_git_sha", lambda: "abc1234")
    base = run_id_mod.generate_run_id("GPT small")
    assert base == "20310506_070809_abc1234_gpt_small"
    assert run_id_mod.generate_run_id("GPT small", tmp_path) == base
    (tmp_path / base).mkdir()
    assert run_id_mod.generate_run_id("GPT small", tmp_path) == base + "__01"
    for n in range(1, 100):
        (tmp_path / f"{base}__{n:02d}").mkdir()
    with pytest.raises(RuntimeError, match="collision limit"):
        run_id_mod.generate_run_id("GPT small", tmp_path)


def test_short_sha_falls_back_to_nogit(monkeypatch: pytest.MonkeyPatch) -> None:
    def boom(*_a, **_k):  # noqa: ANN002, ANN003, ANN202
        raise subprocess.CalledProcessError(128, ["git"])

    monkeypatch.setattr(run_id_mod.subprocess, "run", boom)
    assert run_id_mod._get_short_git_sha() == "nogit"


def test_run_directory_layout_and_snapshots(tmp_path: Path) -> None:
    cfg = config(tmp_path)
    run_dir = create_run_directory(tmp_path / "runs", "r1")
    assert (run_dir / "logs").is_dir()
    with pytest.raises(FileExistsError):
        create_run_directory(tmp_path / "runs", "r1")
    snap = write_resolved_config(run_dir, cfg)
    assert yaml.safe_load(snap.read_text()) == cfg.model_dump()
    assert not (run_dir / "config.yaml.tmp").exists()
    meta = generate_meta(run_id="r1", run_name="unit", config_path="c.yaml",
                         resolved_config_path="/abs/c.yaml")
    path = write_meta_json(run_dir, meta)
    loaded = json.loads(path.read_text())
    assert loaded["meta_version"] == 1 and loaded["run_id"] == "r1"
    assert loaded["created_at"].endswith("Z") and set(loaded["ddp_env"]) == {
        "RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"}
    assert {"git_sha", "python_version", "platform", "argv", "cwd", "hostname", "pid"} <= set(loaded)


def test_json_log_records_and_handler_idempotence(tmp_path: Path) -> None:
    record = logging.LogRecord("llmtrain.x", logging.INFO, __file__, 1, "hello %s", ("w",), None)
    doc = json.loads(JsonFormatter().format(record))
    assert doc["message"] == "hello w" and doc["level"] == "INFO" and doc["logger"] == "llmtrain.x"
    assert "timestamp" in doc

    name = "llmtrain.test_logger"
    log_file = tmp_path / "logs" / "t.log"
    first = configure_logging(name=name, log_to_file=True, file_name=str(log_file))
    second = configure_logging(name=name, log_to_file=True, file_name=str(log_file))
    assert first is second and len(second.handlers) == 2 and second.propagate is False
    second.info("to file")
    assert "to file" in log_file.read_text()
    third = configure_logging(name=name, log_to_file=False)
    assert len(third.handlers) == 1


def test_summary_json_and_text(tmp_path: Path, monkeypatch: pytest.MonkeyPatch) -> None:
    monkeypatch.setenv("RANK", "0")
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    cfg = config(tmp_path)
    result = TrainResult(final_step=5, final_loss=1.234, final_val_loss=2.5, total_time=5.678,
                         peak_memory=0.0, val_metrics={"val/loss": 2.5}, first_step_loss=3.0,
                         resumed_from_step=2, parameter_count=10, trainable_parameter_count=9)
    doc = format_run_summary(config=cfg, run_id="r", run_dir=tmp_path, json_output=True,
                             train_result=result, resumed_from="r0")
    assert doc["run_id"] == "r" and doc["ddp"]["env"]["RANK"] == "0"
    assert doc["ddp"]["env"]["WORLD_SIZE"] is None and doc["resumed_from"] == "r0"
    assert doc["training"] == {
        "final_step": 5, "final_loss": 1.234, "first_step_loss": 3.0, "total_time": 5.678,
        "peak_memory": 0.0, "parameter_count": 10, "trainable_parameter_count": 9,
        "final_val_loss": 2.5, "val_metrics": {"val/loss": 2.5}, "resumed_from_step": 2}
    json.dumps(doc)

    text = format_run_summary(config=cfg, run_id="r", run_dir=tmp_path, train_result=result,
                              resumed_from="r0")
    assert text.startswith("Planned run:\n  Run ID: r\n")
    assert "final_step=5 final_loss=1.2340 total_time=5.68s" in text
    assert "Validation: val/loss=2.5000" in text and "Resumed from: r0" in text
    assert "env=[RANK=0, WORLD_SIZE=unset" in text

    dry = format_run_summary(config=cfg, run_id="r", run_dir=tmp_path,
                             resolved_model_adapter="dummy_gpt", resolved_data_module="dummy_text",
                             dry_run_steps_executed=4)
    assert "Dry run: resolved_model_adapter=dummy_gpt resolved_data_module=dummy_text steps_executed=4" in dry
    assert "Training:" not in dry
