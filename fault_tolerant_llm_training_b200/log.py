"""Logging.  Same format string and stream as the reference (utils.py:21-29) so
that users' greps over Slurm output files keep working.  On multi-rank runs only
rank 0 emits INFO records; other ranks emit WARNING and above."""
from __future__ import annotations

import logging
import os

logger = logging.getLogger()  # root logger, as the reference (utils.py:10)

_FORMAT = "%(asctime)s - %(name)s - %(levelname)s - %(message)s"


def init_logger(rank: int | None = None) -> logging.Logger:
    if rank is None:
        rank = int(os.environ.get("RANK", "0"))
    logger.setLevel(logging.INFO)
    for h in list(logger.handlers):
        if getattr(h, "_ftl_handler", False):
            logger.removeHandler(h)
    ch = logging.StreamHandler()
    ch.setLevel(logging.INFO if rank == 0 else logging.WARNING)
    ch.setFormatter(logging.Formatter(_FORMAT))
    ch._ftl_handler = True
    logger.addHandler(ch)
    return logger
