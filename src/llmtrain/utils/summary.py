"""Run summary (printed once by rank 0, as text or JSON).

Field names, ordering and text layout track the reference
(`src/llmtrain/utils/summary.py:19-217`) because downstream tooling greps these lines
(`k8s/test_e2e.sh` looks for ``final_step=``).  The implementation is table-driven: each
section lists the config attributes it echoes, and both renderings are derived from that.
"""

from __future__ import annotations

from pathlib import Path
from typing import TYPE_CHECKING, Any

from llmtrain.config.schemas import RunConfig
from llmtrain.utils.metadata import ddp_env_snapshot

if TYPE_CHECKING:  # the result type lives with the trainer; only its attributes are used here
    from llmtrain.training.result import TrainResult

# section key -> (text label, attributes in JSON order, attributes in text order)
_MODEL = ("init block_size d_model n_layers n_heads d_ff dropout tie_embeddings vocab_size").split()
_DATA_JSON = (
    "dataset_name dataset_config text_column cache_dir num_workers train_split val_split"
).split()
_TRAINER = (
    "max_steps micro_batch_size grad_accum_steps lr weight_decay warmup_steps max_grad_norm "
    "log_every_steps eval_every_steps save_every_steps"
).split()
_DDP_JSON = (
    "enabled backend init_method timeout_sec find_unused_parameters rank world_size local_rank "
    "master_addr master_port"
).split()
_DDP_TEXT = _DDP_JSON[:5]
_MLFLOW = "enabled tracking_uri experiment run_name log_models".split()


def _pick(section: Any, names: list[str]) -> dict[str, Any]:
    return {name: getattr(section, name) for name in names}


def _kv(pairs: dict[str, Any]) -> str:
    return " ".join(f"{key}={value}" for key, value in pairs.items())


def _training_block(result: TrainResult) -> dict[str, Any]:
    block: dict[str, Any] = {
        "final_step": result.final_step,
        "final_loss": result.final_loss,
        "first_step_loss": result.first_step_loss,
        "total_time": result.total_time,
        "peak_memory": result.peak_memory,
    }
    for optional in ("parameter_count", "trainable_parameter_count", "final_val_loss"):
        value = getattr(result, optional)
        if value is not None:
            block[optional] = value
    if result.val_metrics:
        block["val_metrics"] = result.val_metrics
    if result.resumed_from_step is not None:
        block["resumed_from_step"] = result.resumed_from_step
    return block


def _training_line(result: TrainResult) -> str:
    line = (
        f"Training: final_step={result.final_step} final_loss={result.final_loss:.4f} "
        f"total_time={result.total_time:.2f}s"
    )
    if result.parameter_count is not None:
        line += f" parameter_count={result.parameter_count}"
    if result.trainable_parameter_count is not None:
        line += f" trainable_parameter_count={result.trainable_parameter_count}"
    if result.final_val_loss is not None:
        line += f" final_val_loss={result.final_val_loss:.4f}"
    if result.resumed_from_step is not None:
        line += f" resumed_from_step={result.resumed_from_step}"
    return line


def format_run_summary(
    *,
    config: RunConfig,
    run_id: str,
    run_dir: str | Path,
    json_output: bool = False,
    resolved_model_adapter: str | None = None,
    resolved_data_module: str | None = None,
    dry_run_steps_executed: int | None = None,
    train_result: TrainResult | None = None,
    resumed_from: str | None = None,
) -> str | dict[str, Any]:
    """Describe the run (and its outcome, when known) as a dict or as indented text."""
    out_dir = Path(run_dir)
    env = ddp_env_snapshot()
    model = {"name": config.model.name, **_pick(config.model, _MODEL)}
    data = {"name": config.data.name, **_pick(config.data, _DATA_JSON)}
    trainer = _pick(config.trainer, _TRAINER)
    mlflow = _pick(config.mlflow, _MLFLOW)
    is_dry_run = any(
        item is not None
        for item in (resolved_model_adapter, resolved_data_module, dry_run_steps_executed)
    )

    if json_output:
        doc: dict[str, Any] = {
            "run_id": run_id,
            "output_dir": str(out_dir),
            "model": model,
            "data": data,
            "trainer": trainer,
            "ddp": {**_pick(config.ddp, _DDP_JSON), "env": env},
            "mlflow": mlflow,
        }
        if resolved_model_adapter is not None:
            doc["resolved_model_adapter"] = resolved_model_adapter
        if resolved_data_module is not None:
            doc["resolved_data_module"] = resolved_data_module
        if dry_run_steps_executed is not None:
            doc["dry_run_steps_executed"] = dry_run_steps_executed
        if resumed_from is not None:
            doc["resumed_from"] = resumed_from
        if train_result is not None:
            doc["training"] = _training_block(train_result)
        return doc

    env_text = ", ".join(f"{key}={value or 'unset'}" for key, value in env.items())
    lines = [
        "Planned run:",
        f"  Run ID: {run_id}",
        f"  Output dir: {out_dir}",
        f"  Model: {_kv(model)}",
        f"  Data: {_kv(data)}",
        f"  Trainer: {_kv(trainer)}",
        f"  DDP: {_kv(_pick(config.ddp, _DDP_TEXT))} env=[{env_text}]",
        f"  MLflow: {_kv(mlflow)}",
    ]
    if resumed_from is not None:
        lines.append(f"  Resumed from: {resumed_from}")
    if is_dry_run:
        lines.append(
            "  Dry run: "
            f"resolved_model_adapter={resolved_model_adapter} "
            f"resolved_data_module={resolved_data_module} "
            f"steps_executed={dry_run_steps_executed}"
        )
    if train_result is not None:
        lines.append("  " + _training_line(train_result))
        if train_result.val_metrics:
            metrics = " ".join(
                f"{key}={value:.4f}" for key, value in sorted(train_result.val_metrics.items())
            )
            lines.append(f"  Validation: {metrics}")
    return "\n".join(lines)
