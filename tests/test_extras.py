"""Example integrations, emission simulator, Pinata garbage collection (SURVEY §2.2 Example/*, §2.3 simulate_rewards,
§2.1 pinata_unpin_old_files)."""
import json

import pytest

from arbius_b200.chain import examples as ex
from arbius_b200.chain.engine_sim import UNIT, Revert, deploy_local
from arbius_b200.chain.simulate_rewards import simulate
from arbius_b200.miner.pinata_gc import list_old_pins, unpin_old_files
from arbius_b200.protocol import template_bytes

APP, MINER, OTHER = "0x" + "c1" * 20, "0x" + "a1" * 20, "0x" + "a2" * 20


def test_example_integrations_full_cycle():
    eng = deploy_local(engine_balance=597000 * UNIT, funded={APP: 10 * UNIT, MINER: 10 * UNIT, OTHER: 10 * UNIT})
    mid = ex.RegisterModel(eng, APP).register_model(APP, 0, template_bytes("anythingv3"))
    look = ex.Lookup(eng, APP)
    assert look.model_address(mid) == APP
    tid = ex.SubmitTask(eng, APP, mid, b'{"prompt":"x","negative_prompt":""}').submit_task()
    assert eng.tasks[tid].owner == APP and eng.tasks[tid].fee == UNIT // 10
    assert look.task_cid(tid) == eng.generate_ipfs_cid(b'{"prompt":"x","negative_prompt":""}')
    for who in (MINER, OTHER):
        eng.token.approve(who, eng.address, 5 * UNIT)
        eng.validator_deposit(who, who, 5 * UNIT)
    assert look.validator_staked_balance(MINER) == 5 * UNIT
    cid = "0x1220" + "cd" * 32
    sol = ex.SubmitSolution(eng, MINER)
    with pytest.raises(Revert, match="non existent commitment"):
        sol.submit_solution(tid, cid)
    sol.signal_commitment(tid, cid)
    sol.submit_solution(tid, cid)
    assert look.solution_cid(tid) == cid
    with pytest.raises(Revert, match="not enough delay"):
        ex.ClaimSolution(eng, MINER).claim_solution(tid)
    eng.time_travel(200)                                   # OTHER's stake is now older than the vote eligibility window
    ex.SubmitContestation(eng, OTHER).submit_contestation(tid)
    assert look.contestation_validator(tid) == OTHER
    with pytest.raises(Revert, match="has contestation"):
        eng.time_travel(2100)
        ex.ClaimSolution(eng, MINER).claim_solution(tid)
    with pytest.raises(Revert, match="not allowed"):
        ex.VoteOnContestation(eng, OTHER).vote_on_contestation(tid, True)      # already voted by contesting
    eng.time_travel(4000)
    before = eng.token.balance_of(APP)
    ex.FinishContestationVote(eng, APP).finish(tid, 2)     # 1 yea vs 1 nay: the accused wins, fees are paid out
    assert eng.solutions[tid].claimed is False and eng.token.balance_of(APP) == before
    # a second task is retracted after the wait
    tid2 = ex.SubmitTask(eng, APP, mid, b'{"prompt":"y","negative_prompt":""}').submit_task()
    with pytest.raises(Revert, match="did not wait long enough"):
        ex.RetractTask(eng, APP).retract_task(tid2)
    eng.time_travel(10001)
    bal = eng.token.balance_of(APP)
    ex.RetractTask(eng, APP).retract_task(tid2)
    assert eng.token.balance_of(APP) == bal + UNIT // 10 * 9 // 10 and tid2 not in eng.tasks


def test_emission_simulator_follows_the_target_curve():
    rows = simulate(years=2.0, tasks_per_hour=1000.0, samples=8)
    assert rows[0]["reward_per_task"] == 1.0 and rows[0]["treasury_share"] == pytest.approx(0.1)
    year1 = min(rows, key=lambda r: abs(r["t_days"] - 365))
    assert year1["target_supply"] == pytest.approx(300000, rel=0.02)        # half of the 600k emission after one year
    assert 0.9 < year1["supply"] / year1["target_supply"] < 1.15             # the difficulty multiplier tracks the target
    rewards = [r["reward_per_task"] for r in rows]
    assert all(a >= b for a, b in zip(rewards, rewards[1:]))


def test_pinata_gc_lists_pages_and_unpins():
    calls = []
    pages = {0: [{"ipfs_pin_hash": f"Qm{i}"} for i in range(3)], 3: [{"ipfs_pin_hash": "QmLast"}]}

    def http(method, url, jwt):
        calls.append((method, url))
        assert jwt == "tok"
        if method == "GET":
            off = int(url.split("pageOffset=")[1])
            return json.dumps({"rows": pages.get(off, [])}).encode()
        return b"OK"
    cids = list_old_pins("tok", 7200, now=1_700_000_000, http=http, page_limit=3)
    assert cids == ["Qm0", "Qm1", "Qm2", "QmLast"] and "pinEnd=2023-11-14T20:13:20.000Z" in calls[0][1]
    calls.clear()
    n = unpin_old_files("tok", 7200, http=http, now=1_700_000_000, log=lambda s: None)
    assert n == 3 and [c for c in calls if c[0] == "DELETE"][0][1].endswith("/pinning/unpin/Qm0")
    assert unpin_old_files("tok", 7200, dry_run=True, http=http, now=1_700_000_000, log=lambda s: None) == 3
