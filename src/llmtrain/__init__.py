"""llmtrain-b200: Blackwell-native GPT training with the ``llmtrain`` contract surface.

Same public package name, CLI, config schema and checkpoint format as the reference
(`/root/reference/src/llmtrain/__init__.py:1-10`); the numeric substrate underneath is
hand-written sm_100a CUDA (see ``csrc/``) instead of rented ATen/DDP.
"""

from importlib import metadata as _metadata

__all__ = ["__version__"]


def _detect_version() -> str:
    for dist_name in ("llmtrain-b200", "local-llm-training-k8s"):
        try:
            return _metadata.version(dist_name)
        except _metadata.PackageNotFoundError:
            continue
    return "1.2.0"


__version__ = _detect_version()
