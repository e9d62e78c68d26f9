"""Rank-tagged logging.

Reproduces the reference's log surface (reference ``train.py:16-29``): one
``basicConfig`` at INFO with a ``[Rank N]`` field that is filled by a filter
reading ``RANK`` from the environment (``?`` when unset).  The filter is
attached to the framework logger only, exactly like the reference, so foreign
loggers that lack the field are not touched.
"""
from __future__ import annotations

import logging
import os

LOG_FORMAT = "%(asctime)s - %(name)s - %(levelname)s - [Rank %(rank)s] %(message)s"

_configured = False


class RankLogFilter(logging.Filter):
    """Injects ``record.rank`` from the ``RANK`` env var (torchrun contract)."""

    def filter(self, record: logging.LogRecord) -> bool:
        record.rank = os.environ.get("RANK", "?")
        return True


class _RankDefault(logging.Filter):
    """Root-handler guard: records from other loggers get a rank too, so the
    shared format string never raises on a missing field."""

    def filter(self, record: logging.LogRecord) -> bool:
        if not hasattr(record, "rank"):
            record.rank = os.environ.get("RANK", "?")
        return True


def configure_logging(level: int = logging.INFO) -> None:
    global _configured
    if _configured:
        return
    logging.basicConfig(level=level, format=LOG_FORMAT)
    for h in logging.getLogger().handlers:
        h.addFilter(_RankDefault())
    _configured = True


def get_logger(name: str = "b200ddp") -> logging.Logger:
    logger = logging.getLogger(name)
    if not any(isinstance(f, RankLogFilter) for f in logger.filters):
        logger.addFilter(RankLogFilter())
    return logger


def is_rank0() -> bool:
    return os.environ.get("RANK", "0") == "0"
