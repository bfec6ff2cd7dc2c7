"""Logger set-up for the ``llmtrain`` logger tree.

Same observable behaviour as `src/llmtrain/utils/logging.py:11-90` in the reference: JSON-lines
records with ``timestamp/level/logger/message[/exc_info]``, one stream handler per target stream,
at most one file handler, repeated calls never duplicate handlers, ``propagate`` switched off.
"""

from __future__ import annotations

import json
import logging
import sys
from datetime import datetime, timezone
from pathlib import Path
from typing import TextIO

_PLAIN_FORMAT = "%(asctime)s %(levelname)s %(name)s: %(message)s"


class JsonFormatter(logging.Formatter):
    """One JSON object per record, ASCII-safe."""

    def format(self, record: logging.LogRecord) -> str:  # noqa: A003
        stamp = datetime.fromtimestamp(record.created, tz=timezone.utc).isoformat()
        doc = {
            "timestamp": stamp,
            "level": record.levelname,
            "logger": record.name,
            "message": record.getMessage(),
        }
        if record.exc_info:
            doc["exc_info"] = self.formatException(record.exc_info)
        return json.dumps(doc, ensure_ascii=True)


def _formatter_for(json_output: bool) -> logging.Formatter:
    return JsonFormatter() if json_output else logging.Formatter(_PLAIN_FORMAT)


def _drop_file_handlers(logger: logging.Logger, keep: logging.Handler | None = None) -> None:
    for handler in [h for h in logger.handlers if isinstance(h, logging.FileHandler)]:
        if handler is keep:
            continue
        logger.removeHandler(handler)
        handler.close()


def configure_logging(
    *,
    level: int = logging.INFO,
    name: str = "llmtrain",
    json_output: bool = True,
    log_to_file: bool = False,
    file_name: str = "train.log",
    stream: TextIO | None = None,
) -> logging.Logger:
    """Idempotently attach a stream handler (and optionally one file handler)."""
    logger = logging.getLogger(name)
    logger.setLevel(level)
    logger.propagate = False
    formatter = _formatter_for(json_output)
    target = stream if stream is not None else sys.stdout

    stream_handler: logging.Handler | None = None
    for handler in logger.handlers:
        # FileHandler subclasses StreamHandler, but its .stream is the file object, never `target`.
        if isinstance(handler, logging.StreamHandler) and getattr(handler, "stream", None) is target:
            stream_handler = handler
            break
    if stream_handler is None:
        stream_handler = logging.StreamHandler(target)
        logger.addHandler(stream_handler)
    stream_handler.setFormatter(formatter)

    if not log_to_file:
        _drop_file_handlers(logger)
        return logger

    destination = Path(file_name)
    destination.parent.mkdir(parents=True, exist_ok=True)
    file_handler: logging.Handler | None = None
    for handler in logger.handlers:
        if isinstance(handler, logging.FileHandler) and Path(handler.baseFilename) == destination:
            file_handler = handler
            break
    if file_handler is None:
        _drop_file_handlers(logger)
        file_handler = logging.FileHandler(destination, encoding="utf-8")
        logger.addHandler(file_handler)
    file_handler.setFormatter(formatter)
    return logger
