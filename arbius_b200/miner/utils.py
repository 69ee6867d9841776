"""Small helpers shared by the miner (reference: miner/src/utils.ts:5-49, miner/src/log.ts:6-27)."""
from __future__ import annotations

import logging
import sys
import time
from typing import Callable, Optional, TypeVar

T = TypeVar("T")
log = logging.getLogger("arbius_b200.miner")


def now() -> int:
    return int(time.time())


def expretry(fn: Callable[[], T], tries: int = 10, base: float = 1.5, sleep: Callable[[float], None] = time.sleep) -> Optional[T]:
    """Exponential backoff; returns None after ``tries`` failures, like the reference (callers check for None)."""
    for retry in range(tries):
        try:
            return fn()
        except Exception as e:  # noqa: BLE001 - the reference retries on anything
            seconds = base ** retry
            log.warning("retry request failed (%s), retrying in %.2f", e, seconds)
            sleep(seconds)
    log.error("retry request failed %d times", tries)
    return None


def initialize_logger(log_path: Optional[str] = None, level: int = logging.INFO) -> logging.Logger:
    """Console logger plus optional append-only file: ``<ms> <LEVEL> <file:line> <message>`` (log.ts:13-26)."""
    log.setLevel(level)
    log.handlers.clear()
    h = logging.StreamHandler(sys.stderr)
    h.setFormatter(logging.Formatter("%(asctime)s %(levelname)s %(filename)s:%(lineno)d %(message)s"))
    log.addHandler(h)
    if log_path:
        fh = logging.FileHandler(log_path, mode="a")
        fh.setFormatter(logging.Formatter("%(created)d%(msecs)03d %(levelname)s %(filename)s:%(lineno)d\t%(message)s"))
        log.addHandler(fh)
    return log
