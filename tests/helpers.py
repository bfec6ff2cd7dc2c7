"""Shared builders for the CPU test-suite."""

from __future__ import annotations

from pathlib import Path
from typing import Any

import yaml

from llmtrain.config.schemas import RunConfig


def payload(root_dir: str | Path = "runs", **sections: dict[str, Any]) -> dict[str, Any]:
    base: dict[str, Any] = {
        "schema_version": 1,
        "run": {"name": "unit"},
        "model": {"name": "dummy_gpt"},
        "data": {"name": "dummy_text"},
        "trainer": {"max_steps": 4, "warmup_steps": 0, "micro_batch_size": 2,
                    "grad_accum_steps": 1, "log_every_steps": 2, "eval_every_steps": 2,
                    "save_every_steps": 2},
        "ddp": {},
        "mlflow": {"enabled": False},
        "logging": {"log_to_file": False},
        "output": {"root_dir": str(root_dir)},
    }
    for name, values in sections.items():
        base[name] = {**base.get(name, {}), **values}
    return base


def config(root_dir: str | Path = "runs", **sections: dict[str, Any]) -> RunConfig:
    return RunConfig.model_validate(payload(root_dir, **sections))


TINY_GPT = {"name": "gpt", "block_size": 8, "d_model": 64, "n_layers": 2, "n_heads": 2,
            "d_ff": 128, "dropout": 0.0, "vocab_size": 16}


def write_yaml(path: Path, doc: dict[str, Any]) -> Path:
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_text(yaml.safe_dump(doc, sort_keys=False), encoding="utf-8")
    return path
