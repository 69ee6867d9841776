"""Bit-exact integer restatement of the PRBMath fixed-point routines EngineV1 uses for emission.

EngineV1.targetTs / diffMul (reference: contract/contracts/EngineV1.sol:443-498) call UD60x18/SD59x18
``div``, ``mul`` and ``exp2``. ``exp2`` walks the 64 fractional bits of a 192.64 fixed-point number and
multiplies by round(2^(2^-i) * 2^64) for every set bit; the constants are regenerated here at import
with 80-digit decimals (they match the PUSH9 immediates in the compiled contract, see tests).
Goldens: contract/test/reward.test.ts:153-179.
"""
from __future__ import annotations

from decimal import Decimal, getcontext
from typing import List

UNIT = 10 ** 18
_MAX_UD_EXP2 = 192 * UNIT


def _gen_constants() -> List[int]:
    getcontext().prec = 80
    two = Decimal(2)
    ln2 = two.ln()
    out = []
    for i in range(1, 65):
        v = (ln2 / (two ** i)).exp() * (two ** 64)
        out.append(int(v.to_integral_value(rounding="ROUND_HALF_EVEN")))
    return out


EXP2_CONSTANTS = _gen_constants()        # index 0 <-> bit 63 (2^-1) ... index 63 <-> bit 0 (2^-64)


def _exp2_192x64(x: int) -> int:
    result = 0x800000000000000000000000000000000000000000000000
    for i in range(64):
        if x & (1 << (63 - i)):
            result = (result * EXP2_CONSTANTS[i]) >> 64
    result *= UNIT
    result >>= 191 - (x >> 64)
    return result


def ud_exp2(x: int) -> int:
    if x >= _MAX_UD_EXP2:
        raise OverflowError("PRBMath_UD60x18_Exp2_InputTooBig")
    return _exp2_192x64((x << 64) // UNIT)


def ud_div(x: int, y: int) -> int:
    return (x * UNIT) // y


def sd_div(x: int, y: int) -> int:
    q = (abs(x) * UNIT) // abs(y)
    return q if (x < 0) == (y < 0) else -q


def sd_mul(x: int, y: int) -> int:
    q = (abs(x) * abs(y)) // UNIT
    return q if (x < 0) == (y < 0) else -q


def sd_exp2(x: int) -> int:
    if x < 0:
        if x < -59_794705707972522261:
            return 0
        return (UNIT * UNIT) // ud_exp2(-x)
    return ud_exp2(x)


STARTING_ENGINE_TOKEN_AMOUNT = 600_000 * UNIT
BASE_TOKEN_STARTING_REWARD = UNIT
YEAR = 60 * 60 * 24 * 365


def target_ts(t: int) -> int:
    if t > 3153600000:
        return STARTING_ENGINE_TOKEN_AMOUNT
    e = ud_exp2(ud_div(t, YEAR))
    return STARTING_ENGINE_TOKEN_AMOUNT - ((STARTING_ENGINE_TOKEN_AMOUNT * UNIT * UNIT) // e // UNIT)


def diff_mul(t: int, ts: int) -> int:
    if not (t > 0 and ts > 0):
        raise ValueError("min vals")
    e = target_ts(t)
    d = sd_div(ts, e)
    if d < 933561438102252700:
        return 100 * UNIT
    one, hundred = UNIT, 100 * UNIT
    c = one + (sd_mul(d - one, hundred) - one)
    if c >= 20 * UNIT:
        return 0
    if c < 0:
        return sd_exp2(abs(c))
    return sd_div(one, sd_exp2(c))


def reward(t: int, ts: int) -> int:
    if ts == 0:
        return BASE_TOKEN_STARTING_REWARD
    return ((STARTING_ENGINE_TOKEN_AMOUNT - ts) * BASE_TOKEN_STARTING_REWARD * diff_mul(t, ts)) \
        // STARTING_ENGINE_TOKEN_AMOUNT // UNIT
