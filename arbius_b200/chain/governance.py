"""Governance side of the protocol as in-process state machines: vote-checkpointing token, Governor (Bravo counting,
4 % quorum, timelock control), TimelockController, the L1 token's gateway registration, and the Uniswap-v3 pool
address derivation.

reference: contract/contracts/GovernorV1.sol:12-184 (settings 6575/6575/1e18, quorum 4 %, proposalsCreated /
descriptionHashes / descriptionCids), TimelockV1.sol:6-18, BaseTokenV1.sol:11-99 (ERC20Votes), L1Token.sol:34-112,
libraries/PoolAddress.sol:1-56. Behaviour follows the OpenZeppelin 4.x contracts those files inherit from.
"""
from __future__ import annotations

import bisect
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

from ..protocol import abi
from ..protocol.cid import onchain_cid
from ..protocol.keccak import keccak256
from .engine_sim import ZERO_ADDR, BaseToken, Revert, _addr, _require

PENDING, ACTIVE, CANCELED, DEFEATED, SUCCEEDED, QUEUED, EXPIRED, EXECUTED = range(8)
STATE_NAMES = ("Pending", "Active", "Canceled", "Defeated", "Succeeded", "Queued", "Expired", "Executed")
AGAINST, FOR, ABSTAIN = 0, 1, 2


class VotesToken(BaseToken):
    """BaseTokenV1 with ERC20Votes: delegation and per-block checkpoints of voting power and total supply."""

    def __init__(self, address: str = "0x" + "b0" * 20, clock: Optional[Callable[[], int]] = None, owner: str = ZERO_ADDR):
        super().__init__(address)
        self.clock = clock or (lambda: 0)
        self.owner = _addr(owner)
        self.l1_address = ZERO_ADDR
        self.delegatees: Dict[str, str] = {}
        self._ckpt: Dict[str, List[Tuple[int, int]]] = {}
        self._supply_ckpt: List[Tuple[int, int]] = []
        self.events: List[tuple] = []

    # -- checkpoints
    @staticmethod
    def _push(arr: List[Tuple[int, int]], block: int, value: int) -> None:
        if arr and arr[-1][0] == block:
            arr[-1] = (block, value)
        else:
            arr.append((block, value))

    @staticmethod
    def _lookup(arr: List[Tuple[int, int]], block: int) -> int:
        i = bisect.bisect_right(arr, (block, 1 << 300))
        return arr[i - 1][1] if i else 0

    def get_votes(self, a: str) -> int:
        arr = self._ckpt.get(_addr(a), [])
        return arr[-1][1] if arr else 0

    def get_past_votes(self, a: str, block: int) -> int:
        _require(block < self.clock(), "ERC20Votes: block not yet mined")
        return self._lookup(self._ckpt.get(_addr(a), []), block)

    def get_past_total_supply(self, block: int) -> int:
        _require(block < self.clock(), "ERC20Votes: block not yet mined")
        return self._lookup(self._supply_ckpt, block)

    def delegates(self, a: str) -> str:
        return self.delegatees.get(_addr(a), ZERO_ADDR)

    def _move(self, src: str, dst: str, amount: int) -> None:
        if src == dst or amount == 0:
            return
        now = self.clock()
        if src != ZERO_ADDR:
            self._push(self._ckpt.setdefault(src, []), now, self.get_votes(src) - amount)
        if dst != ZERO_ADDR:
            self._push(self._ckpt.setdefault(dst, []), now, self.get_votes(dst) + amount)

    def delegate(self, sender: str, to: str) -> None:
        s, t = _addr(sender), _addr(to)
        old = self.delegates(s)
        self.delegatees[s] = t
        self.events.append(("DelegateChanged", {"delegator": s, "fromDelegate": old, "toDelegate": t}))
        self._move(old, t, self.balance_of(s))

    # -- ERC20 hooks
    def mint(self, to: str, amount: int) -> None:
        super().mint(to, amount)
        self._push(self._supply_ckpt, self.clock(), self.total_supply)
        self._move(ZERO_ADDR, self.delegates(to), amount)

    def burn(self, account: str, amount: int) -> None:
        a = _addr(account)
        _require(self.balance_of(a) >= amount, "ERC20: burn amount exceeds balance")
        self.balances[a] -= amount
        self.total_supply -= amount
        self._push(self._supply_ckpt, self.clock(), self.total_supply)
        self._move(self.delegates(a), ZERO_ADDR, amount)

    def bridge_burn(self, sender: str, account: str, amount: int) -> None:
        _require(_addr(sender) == self.l2_gateway, "NOT_GATEWAY")
        self.burn(account, amount)

    def transfer(self, sender: str, to: str, amount: int) -> None:
        super().transfer(sender, to, amount)
        self._move(self.delegates(sender), self.delegates(to), amount)

    def transfer_ownership(self, sender: str, to: str) -> None:
        _require(_addr(sender) == self.owner, "Ownable: caller is not the owner")
        self.owner = _addr(to)


class TimelockSim:
    """TimelockController with one proposer/canceller (the governor) and open execution through the governor."""

    def __init__(self, min_delay: int, clock_time: Callable[[], int], address: str = "0x" + "71" * 20):
        self.address = _addr(address)
        self.min_delay = min_delay
        self.now = clock_time
        self.proposers: set = set()
        self.timestamps: Dict[bytes, int] = {}

    @staticmethod
    def hash_operation_batch(targets, values, payloads, predecessor: bytes, salt: bytes) -> bytes:
        return keccak256(abi.encode(["address[]", "uint256[]", "bytes[]", "bytes32", "bytes32"],
                                    [list(targets), list(values), list(payloads), predecessor, salt]))

    def grant_proposer(self, who: str) -> None:
        self.proposers.add(_addr(who))

    def schedule_batch(self, sender, targets, values, payloads, predecessor, salt, delay) -> bytes:
        _require(_addr(sender) in self.proposers, "AccessControl: account is missing role")
        _require(delay >= self.min_delay, "TimelockController: insufficient delay")
        op = self.hash_operation_batch(targets, values, payloads, predecessor, salt)
        _require(op not in self.timestamps, "TimelockController: operation already scheduled")
        self.timestamps[op] = self.now() + delay
        return op

    def is_operation_ready(self, op: bytes) -> bool:
        ts = self.timestamps.get(op, 0)
        return ts > 1 and ts <= self.now()

    def is_operation_done(self, op: bytes) -> bool:
        return self.timestamps.get(op, 0) == 1

    def is_operation_pending(self, op: bytes) -> bool:
        return self.timestamps.get(op, 0) > 1

    def cancel(self, sender, op: bytes) -> None:
        _require(_addr(sender) in self.proposers, "AccessControl: account is missing role")
        _require(self.is_operation_pending(op), "TimelockController: operation cannot be cancelled")
        del self.timestamps[op]

    def execute_batch(self, call: Callable[[str, str, int, bytes], None], targets, values, payloads, predecessor, salt) -> None:
        op = self.hash_operation_batch(targets, values, payloads, predecessor, salt)
        _require(self.is_operation_ready(op), "TimelockController: operation is not ready")
        for t, v, p in zip(targets, values, payloads):
            call(self.address, t, v, p)                    # msg.sender seen by the target is the timelock
        self.timestamps[op] = 1


@dataclass
class Proposal:
    proposer: str
    vote_start: int
    vote_end: int
    targets: List[str]
    values: List[int]
    calldatas: List[bytes]
    description: str
    executed: bool = False
    canceled: bool = False
    votes: List[int] = field(default_factory=lambda: [0, 0, 0])     # against, for, abstain
    voted: Dict[str, bool] = field(default_factory=dict)
    timelock_id: Optional[bytes] = None


class GovernorSim:
    name = "Governor"
    voting_delay = 6575
    voting_period = 6575
    proposal_threshold = 10 ** 18
    quorum_numerator, quorum_denominator = 4, 100

    def __init__(self, token: VotesToken, timelock: TimelockSim, clock_block: Callable[[], int],
                 call: Callable[[str, str, int, bytes], None], address: str = "0x" + "90" * 20):
        self.address = _addr(address)
        self.token, self.timelock, self.block, self.call = token, timelock, clock_block, call
        self.proposals: Dict[int, Proposal] = {}
        self.proposals_created: List[int] = []
        self.description_hashes: Dict[int, bytes] = {}
        self.description_cids: Dict[int, bytes] = {}
        self.events: List[tuple] = []

    @staticmethod
    def hash_proposal(targets: Sequence[str], values: Sequence[int], calldatas: Sequence[bytes], description_hash: bytes) -> int:
        return int.from_bytes(keccak256(abi.encode(["address[]", "uint256[]", "bytes[]", "bytes32"],
                                                   [list(targets), list(values), list(calldatas), description_hash])), "big")

    def quorum(self, block: int) -> int:
        return self.token.get_past_total_supply(block) * self.quorum_numerator // self.quorum_denominator

    def state(self, pid: int) -> int:
        p = self.proposals.get(pid)
        _require(p is not None, "Governor: unknown proposal id")
        if p.executed:
            return EXECUTED
        if p.canceled:
            return CANCELED
        now = self.block()
        if p.vote_start >= now:
            return PENDING
        if p.vote_end >= now:
            return ACTIVE
        against, for_, abstain = p.votes
        if not (self.quorum(p.vote_start) <= for_ + abstain and for_ > against):
            return DEFEATED
        if p.timelock_id is None:
            return SUCCEEDED
        if self.timelock.is_operation_done(p.timelock_id):
            return EXECUTED
        if self.timelock.is_operation_pending(p.timelock_id):
            return QUEUED
        return CANCELED

    def propose(self, sender: str, targets, values, calldatas, description: str) -> int:
        s = _addr(sender)
        _require(self.token.get_past_votes(s, self.block() - 1) >= self.proposal_threshold,
                 "Governor: proposer votes below proposal threshold")
        dh = keccak256(description.encode())
        pid = self.hash_proposal(targets, values, calldatas, dh)
        _require(len(targets) == len(values) == len(calldatas), "Governor: invalid proposal length")
        _require(len(targets) > 0, "Governor: empty proposal")
        _require(pid not in self.proposals, "Governor: proposal already exists")
        start = self.block() + self.voting_delay
        self.proposals[pid] = Proposal(s, start, start + self.voting_period, [_addr(t) for t in targets], list(values),
                                       [bytes(c) for c in calldatas], description)
        self.proposals_created.append(pid)
        self.description_hashes[pid] = dh
        self.description_cids[pid] = onchain_cid(description.encode())
        self.events.append(("ProposalCreated", {"proposalId": pid, "proposer": s, "targets": list(targets), "values": list(values),
                                                "signatures": [""] * len(targets), "calldatas": [bytes(c) for c in calldatas],
                                                "startBlock": start, "endBlock": start + self.voting_period,
                                                "description": description}))
        return pid

    def cast_vote(self, sender: str, pid: int, support: int, reason: str = "") -> int:
        s = _addr(sender)
        _require(self.state(pid) == ACTIVE, "Governor: vote not currently active")
        p = self.proposals[pid]
        _require(not p.voted.get(s), "GovernorCompatibilityBravo: vote already cast")
        _require(support in (AGAINST, FOR, ABSTAIN), "GovernorCompatibilityBravo: invalid vote type")
        weight = self.token.get_past_votes(s, p.vote_start)
        p.voted[s] = True
        p.votes[support] += weight
        self.events.append(("VoteCast", {"voter": s, "proposalId": pid, "support": support, "weight": weight, "reason": reason}))
        return weight

    def _check(self, targets, values, calldatas, description_hash: bytes) -> Tuple[int, Proposal]:
        pid = self.hash_proposal(targets, values, calldatas, description_hash)
        _require(pid in self.proposals, "Governor: unknown proposal id")
        return pid, self.proposals[pid]

    def queue(self, sender: str, targets, values, calldatas, description_hash: bytes) -> int:
        pid, p = self._check(targets, values, calldatas, description_hash)
        _require(self.state(pid) == SUCCEEDED, "Governor: proposal not successful")
        p.timelock_id = self.timelock.schedule_batch(self.address, targets, values, calldatas, b"\x00" * 32, description_hash,
                                                     self.timelock.min_delay)
        self.events.append(("ProposalQueued", {"proposalId": pid, "eta": self.timelock.timestamps[p.timelock_id]}))
        return pid

    def proposal_eta(self, pid: int) -> int:
        p = self.proposals[pid]
        eta = self.timelock.timestamps.get(p.timelock_id, 0) if p.timelock_id else 0
        return 0 if eta == 1 else eta

    def execute(self, sender: str, targets, values, calldatas, description_hash: bytes) -> int:
        pid, p = self._check(targets, values, calldatas, description_hash)
        _require(self.state(pid) in (SUCCEEDED, QUEUED), "Governor: proposal not successful")
        _require(p.timelock_id is not None, "TimelockController: operation is not ready")
        p.executed = True
        try:
            self.timelock.execute_batch(self.call, targets, values, calldatas, b"\x00" * 32, description_hash)
        except Revert:
            p.executed = False
            raise
        self.events.append(("ProposalExecuted", {"proposalId": pid}))
        return pid

    def cancel(self, sender: str, targets, values, calldatas, description_hash: bytes) -> int:
        pid, p = self._check(targets, values, calldatas, description_hash)
        st = self.state(pid)
        _require(st not in (CANCELED, EXPIRED, EXECUTED), "Governor: proposal not active")
        # Bravo compatibility: the proposer may cancel, anyone may once the proposer drops below the threshold
        s = _addr(sender)
        _require(s == p.proposer or self.token.get_past_votes(p.proposer, self.block() - 1) < self.proposal_threshold,
                 "GovernorBravo: proposer above threshold")
        p.canceled = True
        if p.timelock_id is not None and self.timelock.is_operation_pending(p.timelock_id):
            self.timelock.cancel(self.address, p.timelock_id)
        self.events.append(("ProposalCanceled", {"proposalId": pid}))
        return pid


class L1TokenSim(BaseToken):
    """L1 AIUS: premint to the deployer, custom-gateway registration handshake (reference: L1Token.sol:34-112)."""

    def __init__(self, deployer: str, gateway, router, initial_supply: int, address: str = "0x" + "a1" * 20):
        super().__init__(address)
        self.owner = _addr(deployer)
        self.gateway, self.router = gateway, router
        self._should_register = False
        self.balances[self.owner] = initial_supply * 10 ** 18
        self.total_supply = initial_supply * 10 ** 18

    def is_arbitrum_enabled(self) -> int:
        _require(self._should_register, "NOT_EXPECTED_CALL")
        return 0xB1

    def register_token_on_l2(self, sender: str, l2_token: str, **gas_params) -> None:
        _require(_addr(sender) == self.owner, "Ownable: caller is not the owner")
        prev, self._should_register = self._should_register, True
        try:
            self.gateway.register_token_to_l2(self, l2_token)
            self.router.set_gateway(self, self.gateway)
        finally:
            self._should_register = prev


class GatewayMock:
    """test double for the L1 custom gateway / router pair: checks the 0xb1 handshake like the real gateway"""

    def __init__(self):
        self.l1_to_l2: Dict[str, str] = {}
        self.gateways: Dict[str, object] = {}

    def register_token_to_l2(self, token: L1TokenSim, l2_token: str) -> None:
        _require(token.is_arbitrum_enabled() == 0xB1, "NOT_ARB_ENABLED")
        self.l1_to_l2[token.address] = _addr(l2_token)

    def set_gateway(self, token: L1TokenSim, gateway) -> None:
        _require(token.is_arbitrum_enabled() == 0xB1, "NOT_ARB_ENABLED")
        self.gateways[token.address] = gateway


POOL_INIT_CODE_HASH = bytes.fromhex("e34f199b19b2b4f47f68442619d555527d244f78a3297ea89325f843f87b8b54")


def pool_key(token_a: str, token_b: str, fee: int) -> Tuple[str, str, int]:
    a, b = _addr(token_a), _addr(token_b)
    if int(a, 16) > int(b, 16):
        a, b = b, a
    return a, b, fee


def compute_pool_address(factory: str, key: Tuple[str, str, int]) -> str:
    """CREATE2 address of a Uniswap v3 pool (reference: contract/contracts/libraries/PoolAddress.sol:33-55)"""
    t0, t1, fee = key
    _require(int(t0, 16) < int(t1, 16), "")
    salt = keccak256(abi.encode(["address", "address", "uint24"], [t0, t1, fee]))
    return "0x" + keccak256(b"\xff" + bytes.fromhex(_addr(factory)[2:]) + salt + POOL_INIT_CODE_HASH)[12:].hex()
