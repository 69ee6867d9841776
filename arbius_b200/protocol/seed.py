"""Task id -> generator seed (reference: miner/src/utils.ts:15-19, paper/paper.tex:83)."""
from __future__ import annotations

SEED_MODULUS = 0x1FFFFFFFFFFFF0  # Number.MAX_SAFE_INTEGER - 15


def taskid2seed(taskid: str | int | bytes) -> int:
    if isinstance(taskid, bytes):
        v = int.from_bytes(taskid, "big")
    elif isinstance(taskid, str):
        v = int(taskid, 16) if taskid.startswith(("0x", "0X")) else int(taskid)
    else:
        v = int(taskid)
    return v % SEED_MODULUS
