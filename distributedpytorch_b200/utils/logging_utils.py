"""File logging like the reference (train.py:37-39): ./logs/<train_method>.log, append mode, message-only
format.  In multi-process modes only rank 0 attaches the file handler, so N ranks do not interleave writes
into one file."""
from __future__ import annotations

import logging
import os


def setup_logging(train_method: str, directory: str = "./logs", rank: int = 0) -> logging.Logger:
    logger = logging.getLogger()
    logger.setLevel(logging.INFO)
    for h in list(logger.handlers):
        if getattr(h, "_dpt_handler", False):
            logger.removeHandler(h)
    if rank == 0:
        os.makedirs(directory, exist_ok=True)
        handler: logging.Handler = logging.FileHandler(os.path.join(directory, f"{train_method}.log"), mode="a")
    else:
        handler = logging.NullHandler()
    handler.setFormatter(logging.Formatter("%(message)s"))
    handler._dpt_handler = True  # type: ignore[attr-defined]
    logger.addHandler(handler)
    return logger
