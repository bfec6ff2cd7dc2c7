"""Training runtime: trainer, checkpoints, dry run."""

from llmtrain.training.checkpoint import CheckpointManager, CheckpointPayload
from llmtrain.training.result import TrainResult
from llmtrain.training.trainer import Trainer

__all__ = ["CheckpointManager", "CheckpointPayload", "TrainResult", "Trainer"]
