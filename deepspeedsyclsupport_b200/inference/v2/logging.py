"""Logger of the FastGen engine (parity: inference/v2/logging.py): one process-wide logger, created on first use."""
import logging

from ...utils.logging import LoggerFactory

_INSTANCE = []


def inference_logger(level: int = logging.INFO) -> logging.Logger:
    """`level` only matters for the call that creates the logger."""
    if not _INSTANCE:
        _INSTANCE.append(LoggerFactory.create_logger(name="DS-Inference", level=level))
        _INSTANCE[0].debug("Inference logger created.")
    return _INSTANCE[0]
