"""Strict run-configuration schema.

Contract parity with the reference schema (`src/llmtrain/config/schemas.py:8-186`): nine
required top-level sections, every section frozen / ``extra="forbid"`` / defaults validated,
identical field names and defaults so ``print-config`` output is unchanged.

Additive B200 extensions: ``run.device`` accepts ``"cuda"`` and ``ddp.backend``
accepts ``"nccl"`` (torch DDP over NCCL) and ``"nvl"`` (this repo's NVLink peer-memory bucket
engine).  Everything else B200-specific travels through the ``*.extra`` dicts.
"""

from __future__ import annotations

from typing import Any, Literal

from pydantic import BaseModel, ConfigDict, Field, model_validator


class _StrictSection(BaseModel):
    """Shared policy: immutable, unknown keys rejected, defaults validated."""

    model_config = ConfigDict(extra="forbid", frozen=True, validate_default=True)


def _extras() -> Any:
    return Field(default_factory=dict)


class RunSectionConfig(_StrictSection):
    """Run identity, seed and device."""

    name: str
    seed: int = 1337
    device: Literal["cpu", "mps", "cuda"] = "cpu"
    deterministic: bool = True
    notes: str | None = None


class ModelConfig(_StrictSection):
    """Decoder architecture hyper-parameters."""

    name: str
    init: Literal["random"] = "random"
    block_size: int = Field(default=256, ge=8)
    d_model: int = Field(default=384, ge=64)
    n_layers: int = Field(default=6, ge=1)
    n_heads: int = Field(default=6, ge=1)
    d_ff: int = Field(default=1536, ge=64)
    dropout: float = Field(default=0.1, ge=0.0, lt=1.0)
    tie_embeddings: bool = True
    vocab_size: int | None = None
    extra: dict[str, Any] = _extras()

    @model_validator(mode="after")
    def _dims_are_consistent(self) -> ModelConfig:
        if self.d_model % self.n_heads:
            raise ValueError("d_model must be divisible by n_heads")
        if self.d_ff < self.d_model:
            raise ValueError("d_ff must be greater than or equal to d_model")
        return self


class DataConfig(_StrictSection):
    """Data module selection plus optional Hugging Face dataset coordinates."""

    name: str
    cache_dir: str = ".cache/datasets"
    num_workers: int = Field(default=2, ge=0)
    train_split: str = "train"
    val_split: str = "validation"
    dataset_name: str | None = None
    dataset_config: str | None = None
    text_column: str | None = None
    extra: dict[str, Any] = _extras()


class TrainerConfig(_StrictSection):
    """Step counts, batch geometry, optimiser knobs and cadences."""

    max_steps: int = Field(default=1000, ge=1)
    micro_batch_size: int = Field(default=8, ge=1)
    grad_accum_steps: int = Field(default=4, ge=1)
    lr: float = Field(default=3e-4, gt=0.0)
    weight_decay: float = Field(default=0.1, ge=0.0)
    warmup_steps: int = Field(default=100, ge=0)
    max_grad_norm: float = Field(default=1.0, gt=0.0)
    log_every_steps: int = Field(default=10, ge=1)
    eval_every_steps: int = Field(default=100, ge=1)
    save_every_steps: int = Field(default=500, ge=1)
    extra: dict[str, Any] = _extras()

    @model_validator(mode="after")
    def _warmup_fits(self) -> TrainerConfig:
        if self.warmup_steps > self.max_steps:
            raise ValueError("warmup_steps cannot exceed max_steps")
        return self


class DDPConfig(_StrictSection):
    """Data-parallel launch hints; environment variables win over these at runtime."""

    enabled: bool = False
    backend: Literal["gloo", "nccl", "nvl"] = "gloo"
    init_method: Literal["env://"] = "env://"
    timeout_sec: int = Field(default=1800, ge=1)
    find_unused_parameters: bool = False
    rank: int | None = None
    world_size: int | None = None
    local_rank: int | None = None
    master_addr: str | None = None
    master_port: int | None = None


class MLflowConfig(_StrictSection):
    """Experiment tracking (rank 0 only)."""

    enabled: bool = True
    tracking_uri: str = "file:./mlruns"
    experiment: str = "llm-train-k8s"
    run_name: str | None = None
    log_models: bool = False


class LoggingConfig(_StrictSection):
    """Structured logging destination and verbosity."""

    level: Literal["DEBUG", "INFO", "WARNING", "ERROR"] = "INFO"
    json_output: bool = True
    log_to_file: bool = True
    file_name: str = "train.log"


class OutputConfig(_StrictSection):
    """Where run directories live and which snapshots get written."""

    root_dir: str = "runs"
    run_id: str | None = None
    save_config_copy: bool = True
    save_meta_json: bool = True


class RunConfig(_StrictSection):
    """One executable run: all sections are mandatory (they may be empty mappings)."""

    schema_version: int = Field(default=1, ge=1)
    run: RunSectionConfig
    model: ModelConfig
    data: DataConfig
    trainer: TrainerConfig
    ddp: DDPConfig
    mlflow: MLflowConfig
    logging: LoggingConfig
    output: OutputConfig
