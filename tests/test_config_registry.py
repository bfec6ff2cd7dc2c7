"""Schema strictness, loader errors, registries."""

from __future__ import annotations

from pathlib import Path

import pytest
from pydantic import ValidationError

from llmtrain.config.loader import ConfigLoadError, load_and_validate_config
from llmtrain.config.schemas import RunConfig
from tests.helpers import payload, write_yaml


def test_defaults_match_the_reference_contract() -> None:
    cfg = RunConfig.model_validate(payload())
    assert cfg.run.seed == 1337 and cfg.run.device == "cpu" and cfg.run.deterministic
    assert (cfg.model.block_size, cfg.model.d_model, cfg.model.n_layers) == (256, 384, 6)
    assert cfg.model.extra == {} and cfg.data.extra == {} and cfg.trainer.extra == {}
    assert cfg.ddp.backend == "gloo" and cfg.ddp.timeout_sec == 1800
    assert cfg.mlflow.experiment == "llm-train-k8s" and cfg.output.root_dir == "runs"
    assert list(cfg.model_dump()) == ["schema_version", "run", "model", "data", "trainer", "ddp",
                                      "mlflow", "logging", "output"]


@pytest.mark.parametrize(
    "section,values",
    [
        ("model", {"d_model": 384, "n_heads": 7}),
        ("model", {"d_model": 128, "d_ff": 64}),
        ("trainer", {"warmup_steps": 10, "max_steps": 5}),
        ("trainer", {"lr": 0.0}),
        ("run", {"device": "tpu"}),
        ("ddp", {"backend": "mpi"}),
        ("model", {"surprise": 1}),
    ],
)
def test_invalid_values_are_rejected(section: str, values: dict) -> None:
    with pytest.raises(ValidationError):
        RunConfig.model_validate(payload(**{section: values}))


def test_b200_extensions_are_accepted_and_sections_are_frozen() -> None:
    cfg = RunConfig.model_validate(
        payload(run={"device": "cuda"}, ddp={"backend": "nvl"},
                trainer={"extra": {"engine": "b200", "keep_last_k": 2}})
    )
    assert cfg.run.device == "cuda" and cfg.ddp.backend == "nvl"
    assert cfg.trainer.extra["keep_last_k"] == 2
    with pytest.raises(ValidationError):
        cfg.run.seed = 3  # type: ignore[misc]


def test_every_top_level_section_is_required() -> None:
    doc = payload()
    del doc["mlflow"]
    with pytest.raises(ValidationError):
        RunConfig.model_validate(doc)


def test_loader_round_trip_and_error_shapes(tmp_path: Path) -> None:
    good = write_yaml(tmp_path / "ok.yaml", payload())
    cfg, given, resolved = load_and_validate_config(str(good))
    assert cfg.run.name == "unit" and given == str(good) and resolved == good.resolve()

    bad = write_yaml(tmp_path / "bad.yaml", payload(model={"d_model": 384, "n_heads": 7}))
    with pytest.raises(ConfigLoadError) as info:
        load_and_validate_config(str(bad))
    assert "validation failed" in info.value.message and info.value.errors

    (tmp_path / "list.yaml").write_text("- 1\n- 2\n")
    with pytest.raises(ConfigLoadError, match="mapping"):
        load_and_validate_config(str(tmp_path / "list.yaml"))
    (tmp_path / "broken.yaml").write_text("a: [1\n")
    with pytest.raises(ConfigLoadError, match="YAML parse error"):
        load_and_validate_config(str(tmp_path / "broken.yaml"))
    with pytest.raises(ConfigLoadError, match="unable to read"):
        load_and_validate_config(str(tmp_path / "missing.yaml"))
    with pytest.raises(ConfigLoadError, match="non-empty"):
        load_and_validate_config("  ")
    (tmp_path / "empty.yaml").write_text("")
    with pytest.raises(ConfigLoadError, match="validation failed"):
        load_and_validate_config(str(tmp_path / "empty.yaml"))


def test_registries_register_lookup_and_complain() -> None:
    from llmtrain.data.base import DataModule
    from llmtrain.models.base import ModelAdapter
    from llmtrain.registry import data as dreg
    from llmtrain.registry import initialize_registries
    from llmtrain.registry import models as mreg

    initialize_registries()
    initialize_registries()  # idempotent
    assert {"dummy_gpt", "gpt"} <= set(mreg.available_model_adapters())
    assert {"dummy_text", "hf_text", "synthetic_text"} <= set(dreg.available_data_modules())
    assert mreg.RegistryError is not dreg.RegistryError

    with pytest.raises(mreg.RegistryError, match="Unknown model adapter 'nope'. Available: .*gpt"):
        mreg.get_model_adapter("nope")
    with pytest.raises(dreg.RegistryError, match="Unknown data module 'nope'"):
        dreg.get_data_module("nope")
    with pytest.raises(mreg.RegistryError, match="non-empty"):
        mreg.get_model_adapter("   ")
    with pytest.raises(mreg.RegistryError, match="already registered"):

        @mreg.register_model("gpt")
        class _Dup(ModelAdapter):  # pragma: no cover - never instantiated
            pass

    name = "unit_test_plugin"
    if name not in dreg.available_data_modules():

        @dreg.register_data_module(f" {name} ")
        class _Plugin(DataModule):  # pragma: no cover
            pass

    assert dreg.get_data_module(name).__name__ == "_Plugin"
