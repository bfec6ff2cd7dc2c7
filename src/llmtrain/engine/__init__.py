"""Execution engines (torch oracle path and the B200 native runtime)."""

from llmtrain.engine.base import TrainEngine
from llmtrain.engine.torch_engine import TorchEngine

__all__ = ["TorchEngine", "TrainEngine"]
