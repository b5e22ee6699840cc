"""Python side of the levelled logging (C++ side: csrc/core/logging.h).  Level from `allreduce.loglevel` / `akka.loglevel`
(application.conf:22) or AAR_LOGLEVEL; payload tensors are never formatted (the reference prints whole vectors at INFO,
AllreduceWorker.scala:190)."""
from __future__ import annotations

import logging
import os

_LEVELS = {"DEBUG": logging.DEBUG, "INFO": logging.INFO, "WARNING": logging.WARNING, "WARN": logging.WARNING, "ERROR": logging.ERROR,
           "OFF": logging.CRITICAL + 10}


def get_logger(name: str = "aar") -> logging.Logger:
    log = logging.getLogger(name)
    if not log.handlers:
        h = logging.StreamHandler()
        h.setFormatter(logging.Formatter("[%(asctime)s] [%(levelname)s] %(message)s"))
        log.addHandler(h)
        log.setLevel(_LEVELS.get(os.environ.get("AAR_LOGLEVEL", "WARNING").upper(), logging.WARNING))
        log.propagate = False
    return log


def set_level(level: str) -> None:
    """Applies to both runtimes: the Python logger and the C++ one."""
    get_logger().setLevel(_LEVELS.get(level.upper(), logging.INFO))
    try:
        from .. import _core

        _core.set_log_level(level)
    except Exception:
        pass
