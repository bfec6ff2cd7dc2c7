"""CLI through ``python -m llmtrain`` (subprocess) and wiring checks in-process."""

from __future__ import annotations

import json
import os
import subprocess
import sys
from pathlib import Path
from unittest.mock import Mock

import pytest
import yaml

from llmtrain import cli
from tests.helpers import TINY_GPT, payload, write_yaml

SRC = str(Path(__file__).resolve().parents[1] / "src")


def _run(*args: str, cwd: Path | None = None) -> subprocess.CompletedProcess[str]:
    env = dict(os.environ, PYTHONPATH=SRC + os.pathsep + os.environ.get("PYTHONPATH", ""))
    for var in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(var, None)
    return subprocess.run([sys.executable, "-m", "llmtrain", *args], capture_output=True,
                          text=True, check=False, env=env, cwd=cwd, timeout=300)


def test_help_version_and_validate(tmp_path: Path) -> None:
    assert "usage" in _run("--help").stdout.lower()
    assert _run("--version").stdout.startswith("llmtrain ")
    good = write_yaml(tmp_path / "c.yaml", payload(tmp_path / "runs"))
    ok = _run("validate", "--config", str(good))
    assert ok.returncode == 0 and "succeeded" in ok.stdout
    assert json.loads(_run("validate", "--config", str(good), "--json").stdout) == {"status": "ok"}

    bad = write_yaml(tmp_path / "bad.yaml", payload(model={"d_model": 384, "n_heads": 7}))
    res = _run("validate", "--config", str(bad), "--json")
    assert res.returncode == 2 and res.stdout == ""
    err = json.loads(res.stderr)
    assert err["status"] == "error" and "validation failed" in err["message"] and err["errors"]
    plain = _run("validate", "--config", str(bad))
    assert plain.returncode == 2 and plain.stderr.startswith("Config error:")


def test_print_config_yaml_and_json(tmp_path: Path) -> None:
    path = write_yaml(tmp_path / "c.yaml", payload(tmp_path / "runs"))
    as_json = json.loads(_run("print-config", "--config", str(path), "--json").stdout)
    as_yaml = yaml.safe_load(_run("print-config", "--config", str(path)).stdout)
    assert as_json == as_yaml and as_json["model"]["d_model"] == 384
    assert list(as_yaml)[0] == "schema_version"


def test_train_dry_run_writes_run_dir_but_no_checkpoints(tmp_path: Path) -> None:
    path = write_yaml(tmp_path / "c.yaml", payload(tmp_path / "runs"))
    res = _run("train", "--config", str(path), "--run-id", "dry", "--dry-run", "--json")
    assert res.returncode == 0, res.stderr
    doc = json.loads(res.stdout)
    assert doc["run_id"] == "dry" and doc["dry_run_steps_executed"] == 4
    assert doc["resolved_model_adapter"] == "dummy_gpt" and "training" not in doc
    run_dir = tmp_path / "runs" / "dry"
    assert (run_dir / "config.yaml").exists() and (run_dir / "meta.json").exists()
    assert not (run_dir / "checkpoints").exists()
    assert "Dry-run step 1/4" in res.stderr
    dup = _run("train", "--config", str(path), "--run-id", "dry", "--dry-run")
    assert dup.returncode != 0  # run directories are never reused


def test_train_gpt_then_resume_by_run_id(tmp_path: Path) -> None:
    doc = payload(tmp_path / "runs", model=TINY_GPT,
                  trainer={"max_steps": 6, "save_every_steps": 3, "log_every_steps": 3,
                           "eval_every_steps": 6, "lr": 3e-3},
                  logging={"log_to_file": True})
    path = write_yaml(tmp_path / "c.yaml", doc)
    first = _run("train", "--config", str(path), "--run-id", "one", "--json")
    assert first.returncode == 0, first.stderr
    summary = json.loads(first.stdout)  # exactly one JSON document on stdout
    assert summary["training"]["final_step"] == 6 and summary["training"]["parameter_count"] == 68608
    run_dir = tmp_path / "runs" / "one"
    assert sorted(p.name for p in (run_dir / "checkpoints").iterdir()) == [
        "step_000003.pt", "step_000006.pt"]
    assert (run_dir / "logs" / "train.log").read_text().count("checkpoint: saved step") == 2
    assert first.stderr.count("step=") >= 2  # --json moves the trainer's log lines to stderr

    doc["trainer"]["max_steps"] = 8
    longer = write_yaml(tmp_path / "c2.yaml", doc)
    second = _run("train", "--config", str(longer), "--run-id", "two", "--resume", "one")
    assert second.returncode == 0, second.stderr
    assert "Resumed from: one" in second.stdout and "resumed_from_step=6" in second.stdout
    assert "final_step=8" in second.stdout


def test_unknown_plugin_is_a_config_error(tmp_path: Path) -> None:
    path = write_yaml(tmp_path / "c.yaml", payload(tmp_path / "runs", model={"name": "nope"}))
    res = _run("train", "--config", str(path), "--json")
    assert res.returncode == 2 and "Unknown model adapter 'nope'" in json.loads(res.stderr)["message"]


def test_tracker_fallback_and_always_ended(tmp_path: Path, monkeypatch: pytest.MonkeyPatch) -> None:
    from llmtrain.config.schemas import RunConfig
    from llmtrain.tracking import NullTracker

    cfg = RunConfig.model_validate(payload(tmp_path, mlflow={"enabled": True}))

    def unavailable(**_kw):  # noqa: ANN003, ANN202
        raise RuntimeError("mlflow import failed")

    monkeypatch.setattr(cli, "MLflowTracker", unavailable)
    logger = Mock()
    assert isinstance(cli._create_tracker(cfg, logger), NullTracker)
    logger.warning.assert_called_once()

    tracker = Mock()
    monkeypatch.setattr(cli, "_create_tracker", lambda *_a: tracker)
    failing = Mock()
    failing.return_value.fit.side_effect = RuntimeError("boom")
    monkeypatch.setattr(cli, "Trainer", failing)
    path = write_yaml(tmp_path / "c.yaml", payload(tmp_path / "runs"))
    rc = cli.main(["train", "--config", str(path), "--run-id", "x"])
    assert rc == 1
    tracker.start_run.assert_called_once()
    tracker.end_run.assert_called_once()
