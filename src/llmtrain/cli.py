"""``llmtrain`` command line: ``train`` / ``validate`` / ``print-config``.

Flag set, exit codes (0 ok · 1 training or dry-run failure · 2 config / registry error), the
order of side effects, and what is printed where are those of the reference CLI
(`src/llmtrain/cli.py:39-344`; call stack in SURVEY §3.1): rank 0 alone creates the run
directory, writes ``config.yaml`` / ``meta.json``, owns the tracker, and prints the summary;
with ``--json`` all log output is diverted to stderr so stdout carries exactly one document.
"""

from __future__ import annotations

import argparse
import json
import logging
import sys
from collections.abc import Sequence
from pathlib import Path
from typing import Any, TextIO

import yaml

from llmtrain import __version__
from llmtrain.config.loader import ConfigLoadError, load_and_validate_config
from llmtrain.config.schemas import LoggingConfig, RunConfig
from llmtrain.distributed import DDPState, setup_ddp, teardown_ddp
from llmtrain.registry import initialize_registries
from llmtrain.registry.data import RegistryError as DataRegistryError
from llmtrain.registry.data import get_data_module
from llmtrain.registry.models import RegistryError as ModelRegistryError
from llmtrain.registry.models import get_model_adapter
from llmtrain.tracking import MLflowTracker, NullTracker, Tracker
from llmtrain.training import Trainer
from llmtrain.training.dry_run import run_dry_run
from llmtrain.utils.logging import configure_logging
from llmtrain.utils.metadata import generate_meta, write_meta_json
from llmtrain.utils.run_dir import create_run_directory, write_resolved_config
from llmtrain.utils.run_id import generate_run_id
from llmtrain.utils.summary import format_run_summary

EXIT_OK, EXIT_FAILED, EXIT_BAD_CONFIG = 0, 1, 2
_ARTIFACTS = ("config.yaml", "meta.json")


def _configure_logger(
    config_logging: LoggingConfig,
    *,
    verbose: int,
    log_dir: Path | None = None,
    stream: TextIO | None = None,
) -> logging.Logger:
    """Apply the ``logging`` config section (``-v`` forces DEBUG)."""
    level = logging.DEBUG if verbose > 0 else getattr(logging, config_logging.level, logging.INFO)
    target = config_logging.file_name
    if log_dir is not None:
        target = str(log_dir / target)
    return configure_logging(
        level=level,
        json_output=config_logging.json_output,
        log_to_file=config_logging.log_to_file,
        file_name=target,
        stream=stream,
    )


def _emit_config_error(error: ConfigLoadError, *, json_output: bool) -> None:
    """Report a configuration problem on stderr (JSON document or two plain lines)."""
    if json_output:
        doc = {
            "status": "error",
            "message": error.message,
            "details": error.details,
            "errors": error.errors,
        }
        print(json.dumps(doc, indent=2, default=str), file=sys.stderr)
        return
    print(f"Config error: {error.message}", file=sys.stderr)
    if error.details:
        print(error.details, file=sys.stderr)


def _create_tracker(config: RunConfig, logger: logging.Logger) -> Tracker:
    """MLflow when enabled and importable, otherwise the null tracker."""
    if not config.mlflow.enabled:
        return NullTracker()
    try:
        return MLflowTracker(
            tracking_uri=config.mlflow.tracking_uri,
            experiment=config.mlflow.experiment,
            run_name=config.mlflow.run_name,
        )
    except RuntimeError as exc:
        logger.warning("MLflow unavailable; falling back to NullTracker: %s", exc)
        return NullTracker()


def _log_run_artifacts(tracker: Tracker, run_dir: Path) -> None:
    for name in _ARTIFACTS:
        candidate = run_dir / name
        if candidate.exists():
            tracker.log_artifact(candidate, artifact_path="artifacts")


def _mirror_logger_to_stderr(name: str, template: logging.Logger) -> logging.Logger:
    """Give ``name`` its own stderr handler that formats like ``template`` does."""
    mirrored = logging.getLogger(name)
    mirrored.setLevel(template.level)
    mirrored.handlers.clear()
    handler = logging.StreamHandler(sys.stderr)
    handler.setFormatter(template.handlers[0].formatter if template.handlers else None)
    mirrored.addHandler(handler)
    mirrored.propagate = False
    return mirrored


def build_parser() -> argparse.ArgumentParser:
    """Argument parser for the three sub-commands."""
    parser = argparse.ArgumentParser(
        prog="llmtrain",
        description="Distributed GPT training on Kubernetes / torchrun (B200-native runtime).",
    )
    parser.add_argument(
        "--version",
        action="version",
        version=f"%(prog)s {__version__}",
        help="Show the llmtrain version and exit.",
    )
    parser.add_argument(
        "-v", "--verbose", action="count", default=0, help="Increase log verbosity (repeatable)."
    )

    shared = argparse.ArgumentParser(add_help=False)
    shared.add_argument("--config", required=True, help="Path to the YAML configuration file.")
    shared.add_argument("--run-id", help="Override the run ID for the training command.")
    shared.add_argument(
        "--dry-run", action="store_true", help="Stop after run setup (no training execution)."
    )
    shared.add_argument("--json", action="store_true", help="Emit machine-readable JSON output.")

    commands = parser.add_subparsers(dest="command", required=True)
    train = commands.add_parser("train", parents=[shared], help="Run (or dry-run) training.")
    train.add_argument("--resume", default=None, help="Resume from a run_id or checkpoint path.")
    commands.add_parser("validate", parents=[shared], help="Validate a config file.")
    commands.add_parser(
        "print-config", parents=[shared], help="Print the resolved config with defaults."
    )
    return parser


def _load_or_report(args: argparse.Namespace) -> tuple[RunConfig, str, Path] | None:
    try:
        return load_and_validate_config(args.config)
    except ConfigLoadError as exc:
        _emit_config_error(exc, json_output=args.json)
        return None


def _handle_validate(args: argparse.Namespace) -> int:
    loaded = _load_or_report(args)
    if loaded is None:
        return EXIT_BAD_CONFIG
    _configure_logger(
        loaded[0].logging, verbose=args.verbose, stream=sys.stderr if args.json else None
    )
    print(json.dumps({"status": "ok"}, indent=2) if args.json else "Config validation succeeded.")
    return EXIT_OK


def _handle_print_config(args: argparse.Namespace) -> int:
    loaded = _load_or_report(args)
    if loaded is None:
        return EXIT_BAD_CONFIG
    config = loaded[0]
    _configure_logger(
        config.logging, verbose=args.verbose, stream=sys.stderr if args.json else None
    )
    resolved = config.model_dump()
    if args.json:
        print(json.dumps(resolved, indent=2))
    else:
        print(yaml.safe_dump(resolved, sort_keys=False), end="")
    return EXIT_OK


def _handle_train(args: argparse.Namespace) -> int:
    loaded = _load_or_report(args)
    if loaded is None:
        return EXIT_BAD_CONFIG
    config, raw_path, resolved_path = loaded

    ddp_state: DDPState | None = setup_ddp(config) if config.ddp.enabled else None
    is_main = ddp_state is None or ddp_state.is_main

    root_dir = config.output.root_dir
    run_id = args.run_id or config.output.run_id or generate_run_id(config.run.name, root_dir)
    run_dir = create_run_directory(root_dir, run_id) if is_main else Path(root_dir) / run_id

    logger = _configure_logger(
        config.logging,
        verbose=args.verbose,
        log_dir=(run_dir / "logs") if is_main else None,
        stream=sys.stderr if args.json else None,
    )

    if is_main and config.output.save_config_copy:
        write_resolved_config(run_dir, config)
    if is_main and config.output.save_meta_json:
        write_meta_json(
            run_dir,
            generate_meta(
                run_id=run_id,
                run_name=config.run.name,
                config_path=raw_path,
                resolved_config_path=str(resolved_path),
            ),
        )

    initialize_registries()
    try:
        get_model_adapter(config.model.name)
        get_data_module(config.data.name)
    except (ModelRegistryError, DataRegistryError) as exc:
        _emit_config_error(ConfigLoadError(str(exc)), json_output=args.json)
        return EXIT_BAD_CONFIG

    # A real tracker on rank 0 only: concurrent writers would corrupt e.g. an SQLite store.
    tracker: Tracker = _create_tracker(config, logger) if is_main else NullTracker()
    try:
        tracker.start_run(run_name=config.mlflow.run_name or run_id)
        summary_kwargs: dict[str, Any] = {}

        if args.dry_run:
            dry_logger = (
                _mirror_logger_to_stderr("llmtrain.dry_run", logger) if args.json else logger
            )
            try:
                outcome = run_dry_run(config, logger=dry_logger)
            except Exception as exc:
                print(f"Dry-run failed: {exc}", file=sys.stderr)
                return EXIT_FAILED
            summary_kwargs.update(
                resolved_model_adapter=outcome.resolved_model_adapter,
                resolved_data_module=outcome.resolved_data_module,
                dry_run_steps_executed=outcome.steps_executed,
            )
        else:
            if args.json:
                _mirror_logger_to_stderr("llmtrain.training.trainer", logger)
            resume_from = getattr(args, "resume", None)
            if resume_from is not None:
                logger.info("Resuming from: %s", resume_from)
            try:
                trainer = Trainer(
                    config,
                    run_dir=run_dir if is_main else None,
                    tracker=tracker,
                    ddp_state=ddp_state,
                )
                train_result = trainer.fit(resume_from=resume_from)
            except Exception as exc:
                logger.debug("training failure", exc_info=True)
                print(f"Training failed: {exc}", file=sys.stderr)
                return EXIT_FAILED
            summary_kwargs.update(train_result=train_result, resumed_from=resume_from)

        summary = format_run_summary(
            config=config, run_id=run_id, run_dir=run_dir, json_output=args.json, **summary_kwargs
        )
        if is_main:
            _log_run_artifacts(tracker, run_dir)
            print(json.dumps(summary, indent=2) if args.json else summary)
    finally:
        tracker.end_run()
        if ddp_state is not None:
            teardown_ddp()
    return EXIT_OK


_HANDLERS = {
    "validate": _handle_validate,
    "print-config": _handle_print_config,
    "train": _handle_train,
}


def main(argv: Sequence[str] | None = None) -> int:
    """Parse ``argv`` and dispatch; returns the process exit code."""
    args = build_parser().parse_args(argv)
    handler = _HANDLERS.get(args.command)
    if handler is None:
        print(f"Unknown command: {args.command}", file=sys.stderr)
        return EXIT_FAILED
    return handler(args)
