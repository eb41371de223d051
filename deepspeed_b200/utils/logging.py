"""Rank-aware logging helpers.

Behavioural parity with the reference ``deepspeed/utils/logging.py`` (LoggerFactory :22,
log_dist :113, print_json_dist :156, warning_once) -- re-implemented, single code path.
"""
import functools
import json
import logging
import os
import sys

LOG_LEVELS = {
    "debug": logging.DEBUG,
    "info": logging.INFO,
    "warning": logging.WARNING,
    "error": logging.ERROR,
    "critical": logging.CRITICAL,
}


def _build_logger(name="DeepSpeedB200", level=logging.INFO):
    lg = logging.getLogger(name)
    if lg.handlers:
        return lg
    lg.setLevel(level)
    lg.propagate = False
    h = logging.StreamHandler(stream=sys.stdout)
    h.setFormatter(
        logging.Formatter("[%(asctime)s] [%(levelname)s] [%(filename)s:%(lineno)d:%(funcName)s] %(message)s"))
    lg.addHandler(h)
    return lg


class LoggerFactory:
    """Reference ``utils/logging.py:22``: named stdout loggers with the framework's line format."""

    @staticmethod
    def create_logger(name=None, level=logging.WARNING):
        if name is None:
            raise ValueError("name for logger cannot be None")
        return _build_logger(name, level)


def get_log_level_from_string(log_level_str):
    key = str(log_level_str).lower()
    if key not in LOG_LEVELS:
        raise ValueError(f"{log_level_str} is not a valid log level. Valid log levels are {list(LOG_LEVELS)}.")
    return LOG_LEVELS[key]


def should_log_le(max_log_level_str):
    """True when the framework logger's level is at or below the named one (``should_log_le("info")``: info output
    is on)."""
    if not isinstance(max_log_level_str, str):
        raise ValueError(f"{max_log_level_str} is not a string")
    return logger.getEffectiveLevel() <= get_log_level_from_string(max_log_level_str)


logger = _build_logger(level=LOG_LEVELS.get(os.environ.get("DS_LOG_LEVEL", "info").lower(), logging.INFO))


def set_log_level(level):
    if isinstance(level, str):
        level = LOG_LEVELS[level.lower()]
    logger.setLevel(level)


def _rank():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:
        pass
    return int(os.environ.get("RANK", 0))


def should_log(ranks=None):
    """True when this process should emit (ranks=None/[-1] means every rank)."""
    if ranks is None or -1 in ranks:
        return True
    return _rank() in set(ranks)


def log_dist(message, ranks=None, level=logging.INFO):
    """Log ``message`` only on the listed ranks (default rank 0 when ranks is None -> all)."""
    if should_log(ranks):
        logger.log(level, f"[Rank {_rank()}] {message}")


@functools.lru_cache(None)
def log_dist_once(message, ranks=None, level=logging.INFO):
    """``log_dist`` that emits a given (message, ranks, level) only the first time."""
    log_dist(message, ranks=tuple(ranks) if ranks is not None else None, level=level)


def set_log_level_from_string(log_level_str, custom_logger=None):
    """Reference ``utils/logging.py``: 'debug' | 'info' | 'warning' | 'error' | 'critical'."""
    key = str(log_level_str).lower()
    if key not in LOG_LEVELS:
        raise ValueError(f"{log_level_str} is not one of the valid logging levels. Valid log levels are "
                         f"{list(LOG_LEVELS)}.")
    (custom_logger or logger).setLevel(LOG_LEVELS[key])


def get_current_level():
    return logger.getEffectiveLevel()


def print_json_dist(message, ranks=None, path=None):
    """Dump a dict as JSON from the listed ranks, optionally to ``path``."""
    if should_log(ranks):
        message = dict(message)
        message["rank"] = _rank()
        if path is None:
            print(json.dumps(message))
        else:
            with open(path, "w") as f:
                json.dump(message, f)
                os.fsync(f)


@functools.lru_cache(None)
def warning_once(msg):
    logger.warning(msg)


def get_current_level():
    return logger.getEffectiveLevel()
