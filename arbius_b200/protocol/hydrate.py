"""Input hydration + model filters, bug-compatible with the reference miner.

Validators must agree on which inputs are invalid (invalid => contest), so the
quirks of ``hydrateInput`` (reference: miner/src/models.ts:145-220) are kept:

* ``decimal`` is checked with ``col !== (col|0)`` -> non-integers are rejected
  (models.ts:185-189); defaults are never validated (models.ts:210-212).
* the max bound is written ``row > col.max`` and never fires (models.ts:193-197).
* ``int`` is limited to signed 32 bit by ``(col|0)`` (models.ts:179-183).
* type ``file`` has no check at all.

``check_model_filter`` mirrors models.ts:100-143.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Any, Optional, Sequence


def _js_or0(x: float) -> int:
    """ECMAScript ``x|0`` (ToInt32)."""
    if isinstance(x, bool):
        return int(x)
    if isinstance(x, float) and (math.isnan(x) or math.isinf(x)):
        return 0
    v = int(x) & 0xFFFFFFFF  # int() truncates toward zero like ToInt32
    return v - (1 << 32) if v & 0x80000000 else v


def _is_js_number(v: Any) -> bool:
    return isinstance(v, (int, float)) and not isinstance(v, bool)


@dataclass
class HydrationResult:
    input: dict
    err: bool
    errmsg: str


def hydrate_input(preprocessed: Any, template: dict) -> HydrationResult:
    out: dict = {}

    def e(msg: str) -> HydrationResult:
        return HydrationResult(out, True, msg)

    src = preprocessed if isinstance(preprocessed, dict) else {}
    for row in template["input"]:
        var = row["variable"]
        present = var in src
        col = src.get(var)
        if row.get("required") and not present:
            return e(f"input missing required field ({var})")
        if present:
            t = row["type"]
            if t in ("string", "string_enum"):
                if not isinstance(col, str):
                    return e(f"input wrong type ({var})")
            elif t in ("int", "int_enum", "decimal"):
                if not _is_js_number(col) or col != _js_or0(col):
                    return e(f"input wrong type ({var})")
            if t in ("int", "decimal"):
                # `row > col.max` in the reference is always false: only min is enforced.
                if "min" in row and col < row["min"]:
                    return e(f"input out of bounds ({var})")
            if t in ("string_enum", "int_enum"):
                if col not in row.get("choices", []):
                    return e(f"input not in enum ({var})")
                # JS `includes` is strict: 768.0 === 768 but "768" !== 768 and true !== 1
                if isinstance(col, bool):
                    return e(f"input not in enum ({var})")
            out[var] = col
        else:
            out[var] = row.get("default")
    return HydrationResult(out, False, "")


@dataclass
class MiningFilter:
    minfee: int = 0
    mintime: int = 0
    owner: Optional[str] = None


@dataclass
class FilterResult:
    model_enabled: bool
    filter_passed: bool
    model_template: Optional[dict]


@dataclass
class ModelEntry:
    """Registry row (reference: miner/src/types.ts:86-101)."""
    id: str
    template: dict
    mineable: bool = True
    filters: Sequence[MiningFilter] = field(default_factory=lambda: [MiningFilter()])
    name: str = ""


def get_model_by_id(models: Sequence[ModelEntry], model_id: str) -> Optional[ModelEntry]:
    for m in models:
        if m.id.lower() == model_id.lower():
            return m
    return None


def check_model_filter(models: Sequence[ModelEntry], *, model: str, now: int, fee: int,
                       blocktime: int, owner: str) -> FilterResult:
    m = get_model_by_id(models, model)
    if m is None:
        return FilterResult(False, False, None)
    passed = False
    for f in m.filters:
        if f.owner and owner.lower() != f.owner.lower():
            continue
        if fee < f.minfee:
            continue
        if f.mintime > 0 and (now - blocktime) < f.mintime:
            continue
        passed = True
        break
    return FilterResult(True, passed, m.template)
