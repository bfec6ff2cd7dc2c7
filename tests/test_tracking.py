"""Tracker contract: null tracker, mlflow adapter against a fake module."""

from __future__ import annotations

import sys
from types import ModuleType
from unittest.mock import Mock

import pytest

from llmtrain.tracking import MLflowTracker, NullTracker, Tracker
from llmtrain.tracking.mlflow import _flatten_params


def test_null_tracker_accepts_everything() -> None:
    t = NullTracker()
    assert isinstance(t, Tracker)
    t.start_run("x", run_id=None)
    t.log_params({"a": 1})
    t.log_metrics({"m": 1.0}, step=3)
    t.log_artifact("file", artifact_path="artifacts")
    assert t.end_run() is None


def test_flatten_rules() -> None:
    flat = _flatten_params({"a": {"b": 1, "c": {"d": None}}, "l": [1, "x"], "t": (1, 2), "s": "v",
                            "o": object})
    assert flat["a.b"] == 1 and flat["a.c.d"] == "None" and flat["l"] == '[1, "x"]'
    assert flat["t"] == "[1, 2]" and flat["s"] == "v" and flat["o"].startswith("<class")


def test_mlflow_adapter_call_contract(monkeypatch: pytest.MonkeyPatch) -> None:
    fake = ModuleType("mlflow")
    for name in ("set_tracking_uri", "set_experiment", "start_run", "log_params", "log_metrics",
                 "log_artifact", "end_run"):
        setattr(fake, name, Mock())
    fake.active_run = Mock(return_value=Mock(info=Mock(run_id="abc")))
    monkeypatch.setitem(sys.modules, "mlflow", fake)

    t = MLflowTracker(tracking_uri="sqlite:///x.db", experiment="exp", run_name="default")
    t.start_run()
    fake.set_tracking_uri.assert_called_once_with("sqlite:///x.db")
    fake.set_experiment.assert_called_once_with("exp")
    fake.start_run.assert_called_once_with(run_name="default")
    t.start_run(run_id="join-me")
    fake.start_run.assert_called_with(run_id="join-me")
    assert t.active_run_id == "abc"
    t.log_params({"a": {"b": 1}})
    fake.log_params.assert_called_once_with({"a.b": 1})
    t.log_params({})
    assert fake.log_params.call_count == 1
    t.log_metrics({"x": 1}, step=2)
    fake.log_metrics.assert_called_once_with({"x": 1.0}, step=2)
    t.log_metrics({"x": 1})
    fake.log_metrics.assert_called_with({"x": 1.0})
    t.log_metrics({})
    assert fake.log_metrics.call_count == 2
    t.log_artifact("p", artifact_path="artifacts")
    fake.log_artifact.assert_called_once_with("p", artifact_path="artifacts")
    t.end_run()
    fake.end_run.assert_called_once()


def test_missing_mlflow_raises_runtime_error(monkeypatch: pytest.MonkeyPatch) -> None:
    monkeypatch.setitem(sys.modules, "mlflow", None)
    with pytest.raises(RuntimeError, match="optional 'mlflow' dependency"):
        MLflowTracker(tracking_uri="x", experiment="y")
