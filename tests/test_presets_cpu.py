"""Every shipped preset must load through the strict schema, name registered plugins, and — for the
GPT presets — describe a model the native B200 runtime accepts (head widths 16 / 32 / 64)."""

from __future__ import annotations

from pathlib import Path

import pytest

PRESETS = sorted((Path(__file__).resolve().parents[1] / "configs" / "presets").glob("*.yaml"))


@pytest.mark.parametrize("path", PRESETS, ids=lambda p: p.stem)
def test_preset_validates_and_resolves_plugins(path: Path) -> None:
    from llmtrain.config.loader import load_and_validate_config
    from llmtrain.registry import initialize_registries
    from llmtrain.registry.data import get_data_module
    from llmtrain.registry.models import get_model_adapter

    initialize_registries()
    cfg, given, resolved = load_and_validate_config(str(path))
    assert resolved.name == path.name and given == str(path)
    assert get_model_adapter(cfg.model.name) is not None
    assert get_data_module(cfg.data.name) is not None
    assert cfg.trainer.warmup_steps <= cfg.trainer.max_steps


@pytest.mark.parametrize("path", [p for p in PRESETS if "gpt" in p.stem and "dummy" not in p.stem],
                         ids=lambda p: p.stem)
def test_gpt_presets_fit_the_native_runtime(path: Path) -> None:
    from llmtrain.config.loader import load_and_validate_config
    from llmtrain.engine.b200 import supports
    from llmtrain.models.gpt import GPT

    cfg, _, _ = load_and_validate_config(str(path))
    m = cfg.model
    if m.name != "gpt":
        pytest.skip("not a GPT preset")
    tiny = GPT(vocab_size=16, block_size=8, d_model=m.d_model, n_layers=1, n_heads=m.n_heads,
               d_ff=m.d_ff, dropout=m.dropout)
    assert supports(tiny), f"head_dim {m.d_model // m.n_heads} is outside the native coverage"
