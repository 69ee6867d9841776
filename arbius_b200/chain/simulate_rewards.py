"""Emission simulator: what the reference's unfinished ``contract/simulate_rewards`` stub set out to do
(reference: contract/simulate_rewards/src/run.ts:1-14 — a list of users with balances and nothing else).

Runs the Engine's reward law (EngineV1.sol:443-533: targetTs, diffMul, reward — bit-exact PRBMath in prbmath.py) forward
in time for a given solve rate and reports, per sample, the time, the emitted supply, the target supply, the difficulty
multiplier and the per-task reward, split as claimSolution splits it (treasury share, validator share).

  python -m arbius_b200.chain.simulate_rewards --years 4 --tasks-per-hour 1000 --samples 16
"""
from __future__ import annotations

import argparse
import json
from typing import List

from . import prbmath

UNIT = 10 ** 18
STARTING_ENGINE_TOKEN_AMOUNT = 600_000 * UNIT


def simulate(years: float = 4.0, tasks_per_hour: float = 1000.0, samples: int = 16, rate: int = UNIT,
             treasury_reward_percentage: int = 10 ** 17) -> List[dict]:
    """Step the emission forward in blocks of time; every solved task pays reward(t, supply) * rate."""
    horizon = int(years * 365 * 86400)
    step = max(1, horizon // max(samples * 64, 1))
    tasks_per_step = tasks_per_hour * step / 3600.0
    supply, out, next_sample = 0, [], 0
    for t in range(0, horizon + 1, step):
        r = prbmath.reward(t, supply) if supply < STARTING_ENGINE_TOKEN_AMOUNT else 0
        per_task = r * rate // UNIT
        if t >= next_sample:
            target = prbmath.target_ts(t)
            treasury = per_task - per_task * (UNIT - treasury_reward_percentage) // UNIT
            out.append({"t_days": round(t / 86400, 2), "supply": supply / UNIT, "target_supply": target / UNIT,
                        "diff_mul": prbmath.diff_mul(t, supply) / UNIT if supply > 0 else None,
                        "reward_per_task": per_task / UNIT, "validator_share": (per_task - treasury) / UNIT,
                        "treasury_share": treasury / UNIT})
            next_sample += horizon // max(samples, 1)
        supply = min(STARTING_ENGINE_TOKEN_AMOUNT, supply + int(per_task * tasks_per_step))
    return out


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--years", type=float, default=4.0)
    ap.add_argument("--tasks-per-hour", type=float, default=1000.0)
    ap.add_argument("--samples", type=int, default=16)
    a = ap.parse_args(argv)
    for row in simulate(a.years, a.tasks_per_hour, a.samples):
        print(json.dumps(row))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
