"""EngineSim vs the goldens of the reference's Hardhat suites (contract/test/base.test.ts, reward.test.ts)."""
import pytest

from arbius_b200.chain import prbmath
from arbius_b200.chain.engine_sim import UNIT, EngineSim, Revert, deploy_local
from arbius_b200.protocol import generate_commitment

E = UNIT
TESTBUF = bytes.fromhex("746573740a")
TESTCID = "0x1220f4ad8a3bd3189da2ad909ee41148d6893d8c629c410f7f2c7e3fae75aade79c8"
OWNER, USER, MODEL = "0x" + "0a" * 20, "0x" + "11" * 20, "0x" + "33" * 20
V = ["0x" + f"{i:02x}" * 20 for i in (0xa1, 0xa2, 0xa3, 0xa4)]


def ether(x: str) -> int:
    whole, _, frac = x.partition(".")
    return int(whole) * E + int((frac + "0" * 18)[:18])


def _engine(balance: str, funded=None) -> EngineSim:
    return deploy_local(engine_balance=ether(balance), funded=funded or {}, owner=OWNER)


def _validators(eng: EngineSim, addrs, amount="2.4"):
    for a in addrs:
        eng.token.mint(a, ether(amount))
        eng.token.approve(a, eng.address, ether(amount))
        eng.validator_deposit(a, a, ether(amount))


def _solve(eng, validator, taskid, cid=TESTCID):
    eng.signal_commitment(validator, generate_commitment(validator, taskid, cid))
    eng.submit_solution(validator, taskid, cid)


def test_prbmath_goldens():
    # contract/test/reward.test.ts:153-179
    assert prbmath.target_ts(15768000) == 175735931288071485118987
    assert prbmath.target_ts(31536000) == 300000 * E
    assert prbmath.target_ts(315360000) == 599414062500000000000000
    assert prbmath.target_ts(31536000000) == 600000 * E
    assert prbmath.diff_mul(31536000, 250000 * E) == 100 * E
    assert prbmath.diff_mul(31536000, 300000 * E) == E
    assert prbmath.diff_mul(31536000, 305000 * E) == 314980262473718305
    assert prbmath.diff_mul(31536000, 355000 * E) == 3027727226196
    assert prbmath.diff_mul(31536000, 360000 * E) == 0
    assert prbmath.EXP2_CONSTANTS[0] == 0x16A09E667F3BCC909 and prbmath.EXP2_CONSTANTS[1] == 0x1306FE0A31B7152DF


def test_register_model_and_cid():
    eng = _engine("599990")
    mid = eng.register_model(USER, USER, 0, TESTBUF)
    assert eng.models[mid].cid == TESTCID and eng.models[mid].rate == 0
    with pytest.raises(Revert, match="model already registered"):
        eng.register_model(USER, USER, 0, TESTBUF)
    with pytest.raises(Revert, match="address must be non-zero"):
        eng.register_model(USER, "0x" + "00" * 20, 0, TESTBUF)


def test_lifecycle_fee_split_3_09_01():
    """base.test.ts:900-982: model fee 3, task fee 4 -> model 3, validator 0.9, accrued 0.1."""
    eng = _engine("599990")
    _validators(eng, V[:1])
    mid = eng.register_model(USER, MODEL, ether("3"), TESTBUF)
    eng.token.mint(USER, ether("4"))
    eng.token.approve(USER, eng.address, ether("4"))
    with pytest.raises(Revert, match="lower fee than model fee"):
        eng.submit_task(USER, 0, USER, mid, ether("2"), TESTBUF)
    tid = eng.submit_task(USER, 0, USER, mid, ether("4"), bytes.fromhex(TESTCID[2:]))
    assert eng.events[-1].name == "TaskSubmitted" and eng.events[-1].args["id"] == tid
    with pytest.raises(Revert, match="non existent commitment"):
        eng.submit_solution(V[0], tid, TESTCID)
    _solve(eng, V[0], tid)
    with pytest.raises(Revert, match="solution already submitted"):
        eng.submit_solution(V[0], tid, TESTCID)
    with pytest.raises(Revert, match="not enough delay"):
        eng.claim_solution(V[0], tid)
    eng.time_travel(3600)
    eng.claim_solution(V[0], tid)
    assert eng.token.balance_of(USER) == 0
    assert eng.token.balance_of(MODEL) == ether("3")
    assert eng.token.balance_of(V[0]) == ether("0.9")
    assert eng.accrued_fees == ether("0.1")
    with pytest.raises(Revert, match="already claimed"):
        eng.claim_solution(V[0], tid)
    eng.withdraw_accrued_fees(OWNER)
    assert eng.token.balance_of(eng.treasury) == ether("0.1") and eng.accrued_fees == 0


def test_commitment_rules():
    eng = _engine("599990")
    _validators(eng, V[:2])
    mid = eng.register_model(USER, USER, 0, TESTBUF)
    tid = eng.submit_task(USER, 0, USER, mid, 0, TESTBUF)
    c = generate_commitment(V[0], tid, TESTCID)
    eng.signal_commitment(USER, c)                       # anyone may register it
    with pytest.raises(Revert, match="commitment exists"):
        eng.signal_commitment(V[0], c)
    with pytest.raises(Revert, match="non existent commitment"):
        eng.submit_solution(V[1], tid, TESTCID)          # commitment is bound to validator 0
    eng.submit_solution(V[0], tid, TESTCID)
    assert eng.solutions[tid].validator == V[0]


def test_validator_minimum_and_only_validator():
    eng = _engine("597000")                              # supply 3000 -> minimum 2.4
    assert eng.get_validator_minimum() == ether("2.4") and eng.get_slash_amount() == ether("0.3")
    mid = eng.register_model(USER, USER, 0, TESTBUF)
    tid = eng.submit_task(USER, 0, USER, mid, 0, TESTBUF)
    eng.signal_commitment(V[0], generate_commitment(V[0], tid, TESTCID))
    with pytest.raises(Revert, match="min staked too low"):
        eng.submit_solution(V[0], tid, TESTCID)


def test_retract_task():
    eng = _engine("599990")
    mid = eng.register_model(USER, USER, 0, TESTBUF)
    eng.token.mint(USER, ether("1"))
    eng.token.approve(USER, eng.address, ether("1"))
    tid = eng.submit_task(USER, 0, USER, mid, ether("1"), TESTBUF)
    with pytest.raises(Revert, match="did not wait long enough"):
        eng.retract_task(USER, tid)
    with pytest.raises(Revert, match="not owner"):
        eng.retract_task(V[0], tid)
    eng.time_travel(10001)
    eng.retract_task(USER, tid)
    assert eng.token.balance_of(USER) == ether("0.9") and eng.accrued_fees == ether("0.1") and tid not in eng.tasks


def test_failed_contestation_three_voters_payouts():
    """base.test.ts:2680-2764: slash 0.29904, accused +0.14952, other nays +0.07476 each."""
    eng = _engine("597000")
    mid = eng.register_model(USER, USER, 0, TESTBUF)
    tid = eng.submit_task(USER, 0, USER, mid, 0, TESTBUF)
    eng.set_solution_mineable_rate(OWNER, mid, E)
    _validators(eng, V)
    _solve(eng, V[0], tid)
    eng.submit_contestation(V[1], tid)
    eng.vote_on_contestation(V[2], tid, False)
    eng.vote_on_contestation(V[3], tid, False)
    for a in V:
        assert eng.validator(a).staked == ether("2.10096") and eng.token.balance_of(a) == 0
    with pytest.raises(Revert, match="has contestation"):
        eng.claim_solution(V[0], tid)
    with pytest.raises(Revert, match="voting period not ended"):
        eng.contestation_vote_finish(V[0], tid, 3)
    eng.time_travel(4000)
    eng.contestation_vote_finish(V[0], tid, 3)
    # the failed contestation also pays the (tiny, supply is above target) mining reward to the solver
    assert eng.token.balance_of(V[0]) >= ether("0.14952") and eng.token.balance_of(V[0]) - ether("0.14952") < E // 100
    assert eng.token.balance_of(V[1]) == 0
    assert eng.token.balance_of(V[2]) == ether("0.07476") and eng.token.balance_of(V[3]) == ether("0.07476")
    assert eng.validator(V[0]).staked == ether("2.4") and eng.validator(V[1]).staked == ether("2.10096")
    assert eng.validator(V[2]).staked == ether("2.4") and eng.validator(V[3]).staked == ether("2.4")


def test_successful_contestation_refunds_task_fee():
    eng = _engine("597000")
    mid = eng.register_model(USER, USER, 0, TESTBUF)
    eng.token.mint(USER, ether("1"))
    eng.token.approve(USER, eng.address, ether("1"))
    tid = eng.submit_task(USER, 0, USER, mid, ether("1"), TESTBUF)
    _validators(eng, V[:3])
    _solve(eng, V[0], tid)
    eng.submit_contestation(V[1], tid)
    eng.vote_on_contestation(V[2], tid, True)
    assert eng.validator_can_vote(V[2], tid) == 0x03
    eng.time_travel(4000)
    assert eng.validator_can_vote(V[3], tid) == 0x02
    eng.contestation_vote_finish(USER, tid, 2)
    slash = eng.contestations[tid].slash_amount
    assert eng.token.balance_of(USER) == ether("1")                      # fee refunded
    assert eng.token.balance_of(V[1]) == slash - slash // 2               # originator
    assert eng.token.balance_of(V[2]) == (slash - (slash - slash // 2))   # the other yea
    assert eng.validator(V[0]).staked == ether("2.4") - slash             # accused stays slashed


def test_reward_split_nine_to_one():
    """reward.test.ts:189-231: at ts ~ 0 the reward is ~100 AIUS * rate 0.1 -> 9 / 1 split."""
    eng = deploy_local(engine_balance=600_000 * E - 1, owner=OWNER)
    mid = eng.register_model(USER, USER, 0, TESTBUF)
    tid = eng.submit_task(USER, 0, USER, mid, 0, TESTBUF)
    eng.set_solution_mineable_rate(OWNER, mid, E // 10)
    _solve(eng, V[0], tid)                                # minimum is 0 at this supply
    eng.time_travel(3600)
    eng.claim_solution(V[0], tid)
    assert eng.token.balance_of(V[0]) == ether("8.999999999999999999")
    assert eng.token.balance_of(eng.treasury) == ether("1")


def test_pause_and_admin_guards():
    eng = _engine("599990")
    with pytest.raises(Revert, match="not pauser"):
        eng.set_paused(USER, True)
    eng.set_paused(OWNER, True)
    with pytest.raises(Revert, match="paused"):
        eng.register_model(USER, USER, 0, TESTBUF)
    eng.set_paused(OWNER, False)
    with pytest.raises(Revert, match="caller is not the owner"):
        eng.set_version(USER, 1)
    eng.set_version(OWNER, 1)
    assert eng.version == 1 and eng.events[-1].name == "VersionChanged"
    eng.set_min_claim_solution_time(OWNER, 10)
    assert eng.min_claim_solution_time == 10


def test_validator_withdraw_flow():
    eng = _engine("599990")
    _validators(eng, V[:1])
    n = eng.initiate_validator_withdraw(V[0], ether("1"))
    with pytest.raises(Revert, match="wait longer"):
        eng.validator_withdraw(V[0], n, V[0])
    eng.time_travel(86400)
    eng.validator_withdraw(V[0], n, V[0])
    assert eng.token.balance_of(V[0]) == ether("1") and eng.validator(V[0]).staked == ether("1.4")
    n2 = eng.initiate_validator_withdraw(V[0], ether("1"))
    eng.cancel_validator_withdraw(V[0], n2)
    with pytest.raises(Revert, match="request not exist"):
        eng.validator_withdraw(V[0], n2, V[0])
