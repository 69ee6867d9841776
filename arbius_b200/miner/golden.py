"""Golden CIDs for the boot-time determinism self-test.

The reference miner computes the CID of ``{prompt: "arbius test cat", seed: 1337}`` on its one enabled model at every
start and exits on mismatch (reference: miner/src/index.ts:983-1001; golden hard-coded for the A100 container). Here
every enabled model has a golden, and a golden belongs to one (template, native library version, GPU model) triple —
the three things that may legitimately change the bytes. Goldens live in a JSON file next to the miner database
(``selftest.goldens`` in MiningConfig.json overrides the path; the key is additive and optional):

    {"kandinsky2|lib3|NVIDIA B200": "0x1220...", ...}

``python -m arbius_b200.cli golden --write <MiningConfig.json>`` records them on a box whose output has been checked
against other validators; every later boot recomputes and compares, and a mismatch is fatal (exit code 1): a miner whose
library was mis-built or whose GPU differs would otherwise submit CIDs that get contested and its stake slashed.
"""
from __future__ import annotations

import json
import os
import threading
from pathlib import Path
from typing import Dict, Optional

SELFTEST_PROMPT = "arbius test cat"          # index.ts:985-988
SELFTEST_SEED = 1337


def selftest_input(template: dict) -> dict:
    """The reference's fixed probe, completed with the template's defaults (hydrateInput semantics, models.ts:210-212)
    and a stand-in for ``file`` inputs (RVM) that no default can cover."""
    inp = {}
    for row in template.get("input", []):
        var, typ = row["variable"], row.get("type")
        if var == "prompt":
            inp[var] = SELFTEST_PROMPT
        elif typ == "file":
            inp[var] = "selftest://synthetic"
        elif "default" in row:
            inp[var] = row["default"]
    inp["seed"] = SELFTEST_SEED
    return inp


def environment_key(template_name: str) -> str:
    lib, gpu = "nolib", "nogpu"
    try:
        from ..ops import native
        if native.available():
            lib = f"lib{native.load().ab_version()}"
    except Exception:  # noqa: BLE001 - a box without the library still boots the control plane (cog strategy)
        pass
    try:
        import torch
        if torch.cuda.is_available():
            gpu = torch.cuda.get_device_name(0)
    except Exception:  # noqa: BLE001
        pass
    return f"{template_name}|{lib}|{gpu}"


class GoldenStore:
    def __init__(self, path: str | os.PathLike):
        self.path = Path(path)
        self._lock = threading.Lock()

    def load(self) -> Dict[str, str]:
        if not self.path.exists():
            return {}
        return json.loads(self.path.read_text())

    def get(self, key: str) -> Optional[str]:
        return self.load().get(key)

    def put(self, key: str, cid: str) -> None:
        with self._lock:
            d = self.load()
            d[key] = cid
            tmp = self.path.with_suffix(self.path.suffix + ".tmp")
            tmp.write_text(json.dumps(d, indent=1, sort_keys=True) + "\n")
            os.replace(tmp, self.path)


def default_path(config) -> Path:
    raw = getattr(config, "raw", {}) or {}
    p = (raw.get("selftest") or {}).get("goldens")
    if p:
        return Path(p)
    return Path(config.db_path).resolve().parent / "goldens.json"
