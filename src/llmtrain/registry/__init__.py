"""Plugin registries and their deterministic population.

`initialize_registries` imports the built-in plugin modules (idempotent via ``sys.modules``),
as in the reference (`src/llmtrain/registry/__init__.py:7-20`), plus this repo's
``synthetic_text`` data module.
"""

from __future__ import annotations

from importlib import import_module

MODEL_REGISTRY_MODULES: tuple[str, ...] = (
    "llmtrain.models.dummy_gpt",
    "llmtrain.models.gpt",
)
DATA_REGISTRY_MODULES: tuple[str, ...] = (
    "llmtrain.data.dummy_text",
    "llmtrain.data.hf_text",
    "llmtrain.data.synthetic_text",
)


def initialize_registries() -> None:
    """Import every built-in plugin module so its ``@register_*`` decorators run."""
    for dotted in (*MODEL_REGISTRY_MODULES, *DATA_REGISTRY_MODULES):
        import_module(dotted)
