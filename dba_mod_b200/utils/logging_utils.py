"""Run folder + ``log.txt`` logger.

Parity: helper.py:35-48 (folder ``saved_models/model_{name}_{time}``, logger ``"logger"`` at DEBUG with a
file handler on ``log.txt`` plus a stream handler).  Unlike the reference, handlers are attached once per run
folder and can be detached (``close_logger``) so repeated simulations in one process do not multiply output.
"""
from __future__ import annotations

import datetime
import logging
import os
from typing import List

LOGGER_NAME = "logger"


def now_tag() -> str:
    """main.py:93 ``'%b.%d_%H.%M.%S'``."""
    return datetime.datetime.now().strftime("%b.%d_%H.%M.%S")


def make_run_folder(name: str, current_time: str, root: str = "saved_models") -> str:
    path = os.path.join(root, f"model_{name}_{current_time}")
    os.makedirs(path, exist_ok=True)  # reference: os.mkdir + FileExistsError guard, and needs root pre-created
    return path


def open_logger(folder_path: str, stream: bool = True, level=logging.DEBUG) -> List[logging.Handler]:
    logger = logging.getLogger(LOGGER_NAME)
    handlers: List[logging.Handler] = [logging.FileHandler(os.path.join(folder_path, "log.txt"))]
    if stream:
        handlers.append(logging.StreamHandler())
    for h in handlers:
        logger.addHandler(h)
    logger.setLevel(level)
    logger.propagate = False
    return handlers


def close_logger(handlers: List[logging.Handler]) -> None:
    logger = logging.getLogger(LOGGER_NAME)
    for h in handlers:
        logger.removeHandler(h)
        try:
            h.close()
        except Exception:  # noqa: BLE001
            pass
