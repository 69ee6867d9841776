"""EngineSim — an in-process restatement of the EngineV1 protocol state machine.

The reference tests its contracts on an in-process Hardhat EVM (reference: contract/test/base.test.ts)
and its miner only against live chains; neither Node nor Hardhat exists on the build/GPU boxes
(SURVEY §0.9). EngineSim follows contract/contracts/EngineV1.sol function by function (line ranges in
each docstring) with the same revert strings, the same integer arithmetic and the same event names,
so the full miner flow — submitTask -> signalCommitment -> submitSolution -> claimSolution, plus
contestations and slashing — runs and is testable on a CPU-only box. A thin JSON-RPC facade
(chain/rpc_server.py) serves it to out-of-process miners.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

from ..protocol.abi import generate_commitment, hash_model, hash_task
from ..protocol.cid import onchain_cid
from . import prbmath

UNIT = 10 ** 18
MAX_SUPPLY_BASE_TOKEN = 1_000_000 * UNIT
STARTING_ENGINE_TOKEN_AMOUNT = 600_000 * UNIT
MIN_SUPPLY_FOR_VALIDATOR_DEPOSITS = 1_000 * UNIT
MIN_SUPPLY_FOR_SLASHING = 2_000 * UNIT
ZERO_ADDR = "0x" + "00" * 20
MAX_UINT256 = (1 << 256) - 1


class Revert(Exception):
    """A failed ``require``: ``str(e)`` is the revert reason string of the contract."""


def _require(cond: bool, msg: str) -> None:
    if not cond:
        raise Revert(msg)


def _addr(a: str) -> str:
    return a.lower()


def _hex(b) -> str:
    if isinstance(b, (bytes, bytearray)):
        return "0x" + bytes(b).hex()
    return b.lower()


@dataclass
class Event:
    name: str
    args: dict
    block: int
    txhash: str


class BaseToken:
    """ERC20 core of BaseTokenV1 (reference: contract/contracts/BaseTokenV1.sol:11-99): balances, allowances,
    bridgeMint gated to the gateway. Name "Arbius", symbol AIUS, 18 decimals."""

    name, symbol, decimals = "Arbius", "AIUS", 18

    def __init__(self, address: str = "0x" + "b0" * 20):
        self.address = _addr(address)
        self.balances: Dict[str, int] = {}
        self.allowances: Dict[Tuple[str, str], int] = {}
        self.total_supply = 0
        self.l2_gateway = ZERO_ADDR

    def balance_of(self, a: str) -> int:
        return self.balances.get(_addr(a), 0)

    def allowance(self, owner: str, spender: str) -> int:
        return self.allowances.get((_addr(owner), _addr(spender)), 0)

    def approve(self, sender: str, spender: str, amount: int) -> None:
        self.allowances[(_addr(sender), _addr(spender))] = amount

    def mint(self, to: str, amount: int) -> None:
        _require(self.total_supply + amount <= MAX_SUPPLY_BASE_TOKEN, "max supply")
        self.balances[_addr(to)] = self.balance_of(to) + amount
        self.total_supply += amount

    def bridge_mint(self, sender: str, to: str, amount: int) -> None:
        _require(_addr(sender) == self.l2_gateway, "NOT_GATEWAY")
        self.mint(to, amount)

    def transfer(self, sender: str, to: str, amount: int) -> None:
        s = _addr(sender)
        _require(self.balance_of(s) >= amount, "ERC20: transfer amount exceeds balance")
        self.balances[s] = self.balance_of(s) - amount
        self.balances[_addr(to)] = self.balance_of(to) + amount

    def transfer_from(self, spender: str, owner: str, to: str, amount: int) -> None:
        key = (_addr(owner), _addr(spender))
        allowed = self.allowances.get(key, 0)
        if allowed != MAX_UINT256:
            _require(allowed >= amount, "ERC20: insufficient allowance")
            self.allowances[key] = allowed - amount
        self.transfer(owner, to, amount)


@dataclass
class Model:
    fee: int
    addr: str
    rate: int
    cid: str


@dataclass
class Validator:
    staked: int = 0
    since: int = 0
    addr: str = ZERO_ADDR


@dataclass
class Task:
    model: str
    fee: int
    owner: str
    blocktime: int
    version: int
    cid: str


@dataclass
class Solution:
    validator: str
    blocktime: int
    claimed: bool
    cid: str


@dataclass
class Contestation:
    validator: str
    blocktime: int
    finish_start_index: int
    slash_amount: int


_ZERO32 = "0x" + "00" * 32


class EngineSim:
    """State + methods of EngineV1. Every mutating method takes ``sender`` first (msg.sender)."""

    def __init__(self, token: Optional[BaseToken] = None, treasury: str = "0x" + "7e" * 20,
                 owner: str = "0x" + "0a" * 20, address: str = "0x" + "e0" * 20, start_time: int = 1_700_000_000,
                 chain_id: int = 31337):
        # initialize(): EngineV1.sol:240-260
        self.address = _addr(address)
        self.token = token or BaseToken()
        self.treasury = _addr(treasury)
        self.owner = _addr(owner)
        self.pauser = _addr(owner)
        self.paused = False
        self.chain_id = chain_id
        self.now = start_time
        self.block_number = 1
        self.start_block_time = start_time
        self.version = 0
        self.validator_minimum_percentage = 8 * 10 ** 14        # 0.0008 ether
        self.slash_amount_percentage = 10 ** 14                  # 0.0001 ether
        self.solution_fee_percentage = 10 ** 17
        self.retraction_fee_percentage = 10 ** 17
        self.treasury_reward_percentage = 10 ** 17
        self.min_claim_solution_time = 2000
        self.min_retraction_wait_time = 10000
        self.min_contestation_vote_period_time = 4000
        self.max_contestation_validator_stake_since = 120
        self.exit_validator_min_unlock_time = 86400
        self.accrued_fees = 0
        self.prevhash = _ZERO32
        self.models: Dict[str, Model] = {}
        self.validators: Dict[str, Validator] = {}
        self.tasks: Dict[str, Task] = {}
        self.solutions: Dict[str, Solution] = {}
        self.contestations: Dict[str, Contestation] = {}
        self.commitments: Dict[str, int] = {}
        self.contestation_voted: Dict[Tuple[str, str], bool] = {}
        self.contestation_yeas: Dict[str, List[str]] = {}
        self.contestation_nays: Dict[str, List[str]] = {}
        self.withdraw_requests: Dict[Tuple[str, int], Tuple[int, int]] = {}
        self.withdraw_counts: Dict[str, int] = {}
        self.withdraw_pending: Dict[str, int] = {}
        self.task_inputs: Dict[str, bytes] = {}          # calldata kept so miners can "getTransaction"
        self.task_txs: Dict[str, str] = {}
        self.events: List[Event] = []
        self._listeners: List[list] = []      # [callback, cursor into self.events]
        self._tx = 0

    # ---------------------------------------------------------------- chain plumbing
    def subscribe(self, fn: Callable[[Event], None]) -> None:
        """Listeners see events when ``deliver()`` runs (after the emitting transaction has been mined),
        never from inside a half-applied state change."""
        self._listeners.append([fn, len(self.events)])

    def deliver(self) -> int:
        n = 0
        for sub in list(self._listeners):
            while sub[1] < len(self.events):
                ev = self.events[sub[1]]
                sub[1] += 1
                n += 1
                sub[0](ev)
        return n

    def _emit(self, name: str, **args) -> None:
        self.events.append(Event(name, args, self.block_number, self._txhash()))

    def _txhash(self) -> str:
        return "0x" + self._tx.to_bytes(32, "big").hex()

    def _begin_tx(self) -> str:
        """Every transaction is mined in its own block (Hardhat automine semantics)."""
        self._tx += 1
        self.block_number += 1
        self.now += 1
        return self._txhash()

    def mine(self, blocks: int = 1) -> None:
        self.block_number += blocks

    def time_travel(self, seconds: int) -> None:
        self.now += seconds
        self.block_number += 1

    # ---------------------------------------------------------------- views
    def get_psuedo_total_supply(self) -> int:
        b = self.token.balance_of(self.address)
        return 0 if b >= STARTING_ENGINE_TOKEN_AMOUNT else STARTING_ENGINE_TOKEN_AMOUNT - b

    def get_slash_amount(self) -> int:
        ts = self.get_psuedo_total_supply()
        if ts < MIN_SUPPLY_FOR_SLASHING:
            return 0
        return ts - (ts * (UNIT - self.slash_amount_percentage)) // UNIT

    def get_validator_minimum(self) -> int:
        ts = self.get_psuedo_total_supply()
        if ts < MIN_SUPPLY_FOR_VALIDATOR_DEPOSITS:
            return 0
        return ts - (ts * (UNIT - self.validator_minimum_percentage)) // UNIT

    def get_reward(self) -> int:
        return prbmath.reward(self.now - self.start_block_time, self.get_psuedo_total_supply())

    target_ts = staticmethod(prbmath.target_ts)
    diff_mul = staticmethod(prbmath.diff_mul)
    reward = staticmethod(prbmath.reward)

    @staticmethod
    def generate_ipfs_cid(content: bytes) -> str:
        return "0x" + onchain_cid(content).hex()

    @staticmethod
    def generate_commitment(sender: str, taskid: str, cid: str) -> str:
        return generate_commitment(sender, taskid, cid)

    def validator(self, a: str) -> Validator:
        return self.validators.get(_addr(a), Validator(addr=ZERO_ADDR))

    # ---------------------------------------------------------------- modifiers
    def _not_paused(self) -> None:
        _require(not self.paused, "paused")

    def _only_owner(self, sender: str) -> None:
        _require(_addr(sender) == self.owner, "Ownable: caller is not the owner")

    def _only_validator(self, sender: str) -> None:
        s = _addr(sender)
        _require(self.validator(s).staked - self.withdraw_pending.get(s, 0) >= self.get_validator_minimum(),
                 "min staked too low")

    # ---------------------------------------------------------------- admin (EngineV1.sol:264-383)
    def transfer_ownership(self, sender, to):
        self._only_owner(sender)
        self.owner = _addr(to)

    def transfer_treasury(self, sender, to):
        self._only_owner(sender)
        self.treasury = _addr(to)
        self._emit("TreasuryTransferred", to=_addr(to))

    def transfer_pauser(self, sender, to):
        self._only_owner(sender)
        self.pauser = _addr(to)
        self._emit("PauserTransferred", to=_addr(to))

    def set_paused(self, sender, paused: bool):
        _require(_addr(sender) == self.pauser, "not pauser")
        self.paused = paused
        self._emit("PausedChanged", paused=paused)

    def set_solution_mineable_rate(self, sender, model: str, rate: int):
        self._only_owner(sender)
        _require(model in self.models and self.models[model].addr != ZERO_ADDR, "model does not exist")
        self.models[model].rate = rate
        self._emit("SolutionMineableRateChange", id=model, rate=rate)

    def set_version(self, sender, version: int):
        self._only_owner(sender)
        self.version = version
        self._emit("VersionChanged", version=version)

    def _setter(name: str, event: str):   # noqa: N805 - helper used at class-definition time
        def fn(self, sender, amount: int):
            self._only_owner(sender)
            setattr(self, name, amount)
            self._emit(event, amount=amount)
        return fn

    set_validator_minimum_percentage = _setter("validator_minimum_percentage", "ValidatorMinimumPercentageChanged")
    set_slash_amount_percentage = _setter("slash_amount_percentage", "SlashAmountPercentageChanged")
    set_solution_fee_percentage = _setter("solution_fee_percentage", "SolutionFeePercentageChanged")
    set_retraction_fee_percentage = _setter("retraction_fee_percentage", "RetractionFeePercentageChanged")
    set_treasury_reward_percentage = _setter("treasury_reward_percentage", "TreasuryRewardPercentageChanged")
    set_min_claim_solution_time = _setter("min_claim_solution_time", "MinClaimSolutionTimeChanged")
    set_min_retraction_wait_time = _setter("min_retraction_wait_time", "MinRetractionWaitTimeChanged")
    set_min_contestation_vote_period_time = _setter("min_contestation_vote_period_time",
                                                    "MinContestationVotePeriodTimeChanged")
    set_max_contestation_validator_stake_since = _setter("max_contestation_validator_stake_since",
                                                         "MaxContestationValidatorStakeSinceChanged")
    set_exit_validator_min_unlock_time = _setter("exit_validator_min_unlock_time", "ExitValidatorMinUnlockTimeChanged")
    del _setter

    def withdraw_accrued_fees(self, sender) -> None:
        """EngineV1.sol:546-551."""
        txhash = self._begin_tx()
        self._not_paused()
        self.token.transfer(self.address, self.treasury, self.accrued_fees)
        self.accrued_fees = 0

    # ---------------------------------------------------------------- models (EngineV1.sol:557-575)
    def register_model(self, sender: str, addr: str, fee: int, template: bytes) -> str:
        txhash = self._begin_tx()
        self._not_paused()
        _require(_addr(addr) != ZERO_ADDR, "address must be non-zero")
        cid = self.generate_ipfs_cid(template)
        mid = hash_model(addr, fee, 0, cid, sender)
        _require(mid not in self.models, "model already registered")
        self.models[mid] = Model(fee=fee, addr=_addr(addr), rate=0, cid=cid)
        self._emit("ModelRegistered", id=mid)
        return mid

    # ---------------------------------------------------------------- validators (EngineV1.sol:581-672)
    def validator_deposit(self, sender: str, validator: str, amount: int) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        v_addr = _addr(validator)
        self.token.transfer_from(self.address, sender, self.address, amount)
        mn = self.get_validator_minimum()
        v = self.validator(v_addr)
        since = v.since
        if v.staked <= mn and v.staked + amount >= mn:
            since = self.now
        self.validators[v_addr] = Validator(staked=v.staked + amount, since=since, addr=v_addr)
        self._emit("ValidatorDeposit", addr=_addr(sender), validator=v_addr, amount=amount)

    def initiate_validator_withdraw(self, sender: str, amount: int) -> int:
        txhash = self._begin_tx()
        self._not_paused()
        s = _addr(sender)
        _require(self.validator(s).staked - self.withdraw_pending.get(s, 0) >= amount, "")
        unlock = self.now + self.exit_validator_min_unlock_time
        count = self.withdraw_counts.get(s, 0) + 1
        self.withdraw_counts[s] = count
        self.withdraw_requests[(s, count)] = (unlock, amount)
        self.withdraw_pending[s] = self.withdraw_pending.get(s, 0) + amount
        self._emit("ValidatorWithdrawInitiated", addr=s, count=count, unlockTime=unlock, amount=amount)
        return count

    def cancel_validator_withdraw(self, sender: str, count: int) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        s = _addr(sender)
        req = self.withdraw_requests.get((s, count))
        _require(req is not None and req[0] > 0, "request not exist")
        self.withdraw_pending[s] -= req[1]
        del self.withdraw_requests[(s, count)]
        self._emit("ValidatorWithdrawCancelled", addr=s, count=count)

    def validator_withdraw(self, sender: str, count: int, to: str) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        s = _addr(sender)
        req = self.withdraw_requests.get((s, count))
        _require(req is not None and req[0] > 0, "request not exist")
        _require(self.now >= req[0], "wait longer")
        _require(self.validator(s).staked >= req[1], "stake insufficient")
        self.token.transfer(self.address, to, req[1])
        self.validators[s].staked -= req[1]
        self.withdraw_pending[s] -= req[1]
        del self.withdraw_requests[(s, count)]
        self._emit("ValidatorWithdraw", addr=s, to=_addr(to), count=count, amount=req[1])

    # ---------------------------------------------------------------- tasks (EngineV1.sol:681-736)
    def submit_task(self, sender: str, version: int, owner: str, model: str, fee: int, input_: bytes) -> str:
        txhash = self._begin_tx()
        self._not_paused()
        model = _hex(model)
        _require(model in self.models, "model does not exist")
        _require(fee >= self.models[model].fee, "lower fee than model fee")
        cid = self.generate_ipfs_cid(input_)          # reverts above 65536 bytes, like the contract
        tid = hash_task(model, fee, cid, sender, self.prevhash)
        self._emit("TaskSubmitted", id=tid, model=model, fee=fee, sender=_addr(sender))
        self.token.transfer_from(self.address, sender, self.address, fee)
        self.tasks[tid] = Task(model=model, fee=fee, owner=_addr(owner), blocktime=self.now, version=version, cid=cid)
        self.task_inputs[tid] = bytes(input_)
        self.task_txs[tid] = txhash
        self.prevhash = tid
        return tid

    def retract_task(self, sender: str, taskid: str) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        t = self.tasks.get(taskid)
        _require(t is not None and t.owner == _addr(sender), "not owner")
        _require(taskid not in self.solutions, "has solution")
        _require(self.now - t.blocktime > self.min_retraction_wait_time, "did not wait long enough")
        amount_minus_fee = (t.fee * (UNIT - self.retraction_fee_percentage)) // UNIT
        fee = t.fee - amount_minus_fee
        self.token.transfer(self.address, sender, amount_minus_fee)
        self.accrued_fees += fee
        del self.tasks[taskid]
        self._emit("TaskRetracted", id=taskid)

    # ---------------------------------------------------------------- commit / reveal (EngineV1.sol:764-812)
    def signal_commitment(self, sender: str, commitment: str) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        commitment = _hex(commitment)
        _require(self.commitments.get(commitment, 0) == 0, "commitment exists")
        self.commitments[commitment] = self.block_number
        self._emit("SignalCommitment", addr=_addr(sender), commitment=commitment)

    def signal_support(self, sender: str, model: str, support: bool) -> None:
        txhash = self._begin_tx()
        self._only_validator(sender)
        _require(_hex(model) in self.models, "model does not exist")
        self._emit("SignalSupport", addr=_addr(sender), model=_hex(model), supported=support)

    def submit_solution(self, sender: str, taskid: str, cid: str) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        self._only_validator(sender)
        taskid, cid = _hex(taskid), _hex(cid)
        _require(taskid in self.tasks, "task does not exist")
        _require(taskid not in self.solutions, "solution already submitted")
        c = generate_commitment(sender, taskid, cid)
        _require(self.commitments.get(c, 0) > 0, "non existent commitment")
        _require(self.commitments[c] < self.block_number, "commitment must be in past")
        self.solutions[taskid] = Solution(validator=_addr(sender), blocktime=self.now, claimed=False, cid=cid)
        self._emit("SolutionSubmitted", addr=_addr(sender), task=taskid)

    # ---------------------------------------------------------------- claim (EngineV1.sol:819-889)
    def _claim_solution_fees_and_reward(self, taskid: str) -> None:
        t, s = self.tasks[taskid], self.solutions[taskid]
        m = self.models[t.model]
        model_fee = m.fee if m.fee <= t.fee else 0
        if model_fee > 0:
            self.token.transfer(self.address, m.addr, model_fee)
        remaining = t.fee - model_fee
        treasury_fee = remaining - (remaining * (UNIT - self.solution_fee_percentage)) // UNIT
        self.accrued_fees += treasury_fee
        validator_fee = remaining - treasury_fee
        if validator_fee > 0:
            self.token.transfer(self.address, s.validator, validator_fee)
        if m.rate > 0:
            total = (self.get_reward() * m.rate) // UNIT
            if total > 0:
                treasury_reward = total - (total * (UNIT - self.treasury_reward_percentage)) // UNIT
                self.token.transfer(self.address, s.validator, total - treasury_reward)
                self.token.transfer(self.address, self.treasury, treasury_reward)

    def claim_solution(self, sender: str, taskid: str) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        taskid = _hex(taskid)
        _require(taskid in self.solutions, "solution not found")
        _require(taskid not in self.contestations, "has contestation")
        s = self.solutions[taskid]
        _require(s.blocktime < self.now - self.min_claim_solution_time, "not enough delay")
        _require(not s.claimed, "already claimed")
        s.claimed = True
        self._emit("SolutionClaimed", addr=s.validator, task=taskid)
        self._claim_solution_fees_and_reward(taskid)

    # ---------------------------------------------------------------- contestations (EngineV1.sol:893-1106)
    def submit_contestation(self, sender: str, taskid: str) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        self._only_validator(sender)
        taskid = _hex(taskid)
        _require(taskid in self.solutions, "solution does not exist")
        _require(taskid not in self.contestations, "contestation already exists")
        s = self.solutions[taskid]
        _require(self.now < s.blocktime + self.min_claim_solution_time, "too late")
        _require(not s.claimed, "wtf")
        slash = self.get_slash_amount()
        self.contestations[taskid] = Contestation(validator=_addr(sender), blocktime=self.now, finish_start_index=0,
                                                  slash_amount=slash)
        self._emit("ContestationSubmitted", addr=_addr(sender), task=taskid)
        self._vote(taskid, True, sender)
        if self.validator(s.validator).staked >= slash:
            self._vote(taskid, False, s.validator)

    def validator_can_vote(self, addr: str, taskid: str) -> int:
        a, taskid = _addr(addr), _hex(taskid)
        c = self.contestations.get(taskid)
        if c is None:
            return 0x01
        if self.now > c.blocktime + self.min_contestation_vote_period_time:
            return 0x02
        if self.contestation_voted.get((taskid, a)):
            return 0x03
        v = self.validator(a)
        if v.since == 0:
            return 0x04
        if v.since < self.max_contestation_validator_stake_since:
            return 0x05
        if v.since - self.max_contestation_validator_stake_since > c.blocktime:
            return 0x06
        return 0x00

    def _vote(self, taskid: str, yea: bool, addr: str) -> None:
        a = _addr(addr)
        self.contestation_voted[(taskid, a)] = True
        (self.contestation_yeas if yea else self.contestation_nays).setdefault(taskid, []).append(a)
        v = self.validators.setdefault(a, Validator(addr=a))
        slash = self.contestations[taskid].slash_amount
        _require(v.staked >= slash, "Arithmetic operation underflowed or overflowed outside of an unchecked block")
        v.staked -= slash
        self._emit("ContestationVote", addr=a, task=taskid, yea=yea)

    def vote_on_contestation(self, sender: str, taskid: str, yea: bool) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        self._only_validator(sender)
        taskid = _hex(taskid)
        _require(self.validator_can_vote(sender, taskid) == 0, "not allowed")
        self._vote(taskid, yea, sender)

    def contestation_vote_finish(self, sender: str, taskid: str, amnt: int) -> None:
        txhash = self._begin_tx()
        self._not_paused()
        taskid = _hex(taskid)
        c = self.contestations.get(taskid)
        _require(c is not None, "contestation doesn't exist")
        _require(self.now >= c.blocktime + self.min_contestation_vote_period_time, "voting period not ended")
        _require(amnt > 0, "amnt too small")
        yeas = self.contestation_yeas.get(taskid, [])
        nays = self.contestation_nays.get(taskid, [])
        ny, nn = len(yeas), len(nays)
        start, end = c.finish_start_index, c.finish_start_index + amnt
        slash = c.slash_amount
        if ny > nn:
            total = nn * slash
            to_orig = total if ny == 1 else total - total // 2
            to_others = 0 if ny == 1 else (total - to_orig) // (ny - 1)
            for i in range(start, min(end, ny)):
                a = yeas[i]
                self.validators[a].staked += slash
                self.token.transfer(self.address, a, to_orig if i == 0 else to_others)
            if c.finish_start_index == 0:
                t = self.tasks[taskid]
                self.token.transfer(self.address, t.owner, t.fee)
        else:
            total = ny * slash
            to_accused = total if nn == 1 else total // 2
            to_others = 0 if nn == 1 else (total - to_accused) // (nn - 1)
            for i in range(start, min(end, nn)):
                a = nays[i]
                self.validators[a].staked += slash
                self.token.transfer(self.address, a, to_accused if i == 0 else to_others)
            if c.finish_start_index == 0:
                self._claim_solution_fees_and_reward(taskid)
        c.finish_start_index = end
        self._emit("ContestationVoteFinish", id=taskid, start_idx=start, end_idx=end)


def deploy_local(engine_balance: int = STARTING_ENGINE_TOKEN_AMOUNT - 1, funded: Optional[Dict[str, int]] = None,
                 **kw) -> EngineSim:
    """Token + engine wired the way the reference's deploy scripts leave them
    (reference: contract/scripts/003-deploy-core-basic.ts:7-112; tests: base.test.ts:29-80)."""
    token = BaseToken()
    eng = EngineSim(token=token, **kw)
    if engine_balance:
        token.mint(eng.address, engine_balance)
    for a, amt in (funded or {}).items():
        token.mint(a, amt)
    return eng
