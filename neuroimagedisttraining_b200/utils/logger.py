"""Loggers.  ``logger_config`` reproduces the experiment scripts' file logger (``main_sailentgrads.py:184-192``: format
``%(message)s``, mode ``w+``, path ``LOG/<dataset>/<identity>.log``) so downstream log parsing is unchanged; a JSONL
metrics sink is added next to it.  ``logging_config`` is the root-logger formatter of ``fedml_api/utils/logger.py:7-32``."""
from __future__ import annotations

import json
import logging
import os
import time
from typing import Optional


def logger_config(log_path: str, logging_name: str, console: bool = True) -> logging.Logger:
    os.makedirs(os.path.dirname(os.path.abspath(log_path)), exist_ok=True)
    logger = logging.getLogger(logging_name)
    logger.setLevel(logging.DEBUG)
    logger.handlers.clear()
    handler = logging.FileHandler(log_path, mode="w+", encoding="UTF-8")
    handler.setLevel(logging.DEBUG)
    handler.setFormatter(logging.Formatter("%(message)s"))
    logger.addHandler(handler)
    if console:
        ch = logging.StreamHandler()
        ch.setLevel(logging.DEBUG)
        logger.addHandler(ch)
    logger.propagate = False
    return logger


def logging_config(args=None, process_id: int = 0, level=logging.INFO) -> None:
    for h in logging.root.handlers[:]:
        logging.root.removeHandler(h)
    logging.basicConfig(level=level if process_id == 0 else logging.WARNING,
                        format=str(process_id) + " - %(asctime)s %(filename)s[line:%(lineno)d] %(levelname)s %(message)s",
                        datefmt="%a, %d %b %Y %H:%M:%S")


class JsonlMetrics:
    """One JSON object per line: round metrics, timings, anything serialisable."""

    def __init__(self, path: Optional[str]):
        self.path = path
        if path:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            open(path, "w").close()

    def write(self, **kw) -> None:
        if not self.path:
            return
        kw.setdefault("ts", time.time())
        with open(self.path, "a") as f:
            f.write(json.dumps(kw, default=float) + "\n")
