"""Logging helpers (`logger`, `log_dist`, `print_json_dist`).  Parity: deepspeed/utils/logging.py."""
import functools
import json
import logging
import os
import sys

log_levels = {
    "debug": logging.DEBUG,
    "info": logging.INFO,
    "warning": logging.WARNING,
    "error": logging.ERROR,
    "critical": logging.CRITICAL,
}


def _create_logger(name="DeepSpeedB200", level=logging.INFO):
    lg = logging.getLogger(name)
    lg.setLevel(level)
    lg.propagate = False
    if not lg.handlers:
        h = logging.StreamHandler(stream=sys.stdout)
        h.setFormatter(logging.Formatter("[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d:%(funcName)s] %(message)s"))
        lg.addHandler(h)
    return lg


class LoggerFactory:
    """(parity: deepspeed/utils/logging.py LoggerFactory.create_logger)"""

    @staticmethod
    def create_logger(name=None, level=logging.INFO):
        if name is None:
            raise ValueError("name for logger cannot be None")
        return _create_logger(name, level)


logger = _create_logger(level=log_levels.get(os.environ.get("DSB200_LOG_LEVEL", "info"), logging.INFO))


@functools.lru_cache(None)
def warning_once(msg):
    logger.warning(msg)


logger.warning_once = warning_once


def _rank():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:
        pass
    return int(os.environ.get("RANK", "0"))


def should_log(ranks=None):
    if ranks is None:
        return True
    r = _rank()
    return r in ranks or -1 in ranks


def log_dist(message, ranks=None, level=logging.INFO):
    """Log `message` on the given ranks only (ranks=[-1] -> every rank)."""
    if should_log(ranks):
        logger.log(level, f"[Rank {_rank()}] {message}")


def print_json_dist(message, ranks=None, path=None):
    if should_log(ranks):
        message["rank"] = _rank()
        with open(path, "w") as f:
            json.dump(message, f)
            os.fsync(f)


def get_current_level():
    return logger.getEffectiveLevel()


def should_log_le(max_log_level_str):
    return get_current_level() <= log_levels[max_log_level_str.lower()]
