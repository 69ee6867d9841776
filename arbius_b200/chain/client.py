"""Chain access used by the miner: one interface, two transports.

``ChainClient`` is what the miner calls (the ethers ``arbius`` / ``token`` / ``solver`` contract objects of
reference: miner/src/blockchain.ts:22-75 collapsed into one facade). ``SimChainClient`` binds it to an
in-process EngineSim; ``chain/jsonrpc.py`` binds it to a real node over eth JSON-RPC with locally signed
transactions. Amounts are Python ints (wei), ids/cids/addresses are 0x-hex strings.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional

from .engine_sim import ZERO_ADDR, EngineSim, Event, Revert

GAS_LIMITS = {"submitTask": 2_500_000, "signalCommitment": 450_000, "submitSolution": 500_000,
              "claimSolution": 300_000}    # reference: miner/src/index.ts:483, 621, 636, 737


@dataclass
class TaskView:
    model: str
    fee: int
    owner: str
    blocktime: int
    version: int
    cid: str


@dataclass
class SolutionView:
    validator: str
    blocktime: int
    claimed: bool
    cid: str


@dataclass
class ContestationView:
    validator: str
    blocktime: int
    finish_start_index: int
    slash_amount: int


@dataclass
class Receipt:
    transaction_hash: str
    block_number: int


class ChainError(Exception):
    """A reverted or failed transaction; ``reason`` carries the revert string when known."""

    def __init__(self, reason: str):
        super().__init__(reason)
        self.reason = reason


class ChainClient:
    """Interface. ``address`` is the miner wallet; ``solver_address`` the contract that receives stake."""
    address: str
    solver_address: str

    # views
    def version(self) -> int: raise NotImplementedError
    def now(self) -> int: raise NotImplementedError
    def block_number(self) -> int: raise NotImplementedError
    def eth_balance(self) -> int: raise NotImplementedError
    def tasks(self, taskid: str) -> TaskView: raise NotImplementedError
    def solutions(self, taskid: str) -> SolutionView: raise NotImplementedError
    def contestations(self, taskid: str) -> ContestationView: raise NotImplementedError
    def contestation_voted(self, taskid: str, addr: str) -> bool: raise NotImplementedError
    def validator_staked(self, addr: Optional[str] = None) -> int: raise NotImplementedError
    def get_validator_minimum(self) -> int: raise NotImplementedError
    def token_balance(self, addr: Optional[str] = None) -> int: raise NotImplementedError
    def token_allowance(self, owner: str, spender: str) -> int: raise NotImplementedError
    def task_input_bytes(self, taskid: str, txid: str) -> bytes: raise NotImplementedError
    # transactions
    def token_approve(self, spender: str, amount: int) -> Receipt: raise NotImplementedError
    def validator_deposit(self, validator: str, amount: int) -> Receipt: raise NotImplementedError
    def submit_task(self, version: int, owner: str, model: str, fee: int, input_: bytes) -> str: raise NotImplementedError
    def signal_commitment(self, commitment: str) -> Receipt: raise NotImplementedError
    def submit_solution(self, taskid: str, cid: str) -> Receipt: raise NotImplementedError
    def claim_solution(self, taskid: str) -> Receipt: raise NotImplementedError
    def submit_contestation(self, taskid: str) -> Receipt: raise NotImplementedError
    def vote_on_contestation(self, taskid: str, yea: bool) -> Receipt: raise NotImplementedError
    def contestation_vote_finish(self, taskid: str, amnt: int) -> Receipt: raise NotImplementedError
    # events
    def subscribe(self, handler: Callable[[Event], None]) -> None: raise NotImplementedError
    def poll_events(self) -> int: return 0


class SimChainClient(ChainClient):
    def __init__(self, engine: EngineSim, address: str):
        self.engine = engine
        self.address = address.lower()
        self.solver_address = engine.address

    def _tx(self, fn, *args) -> Receipt:
        try:
            fn(*args)
        except Revert as e:
            raise ChainError(str(e)) from e
        return Receipt(self.engine._txhash(), self.engine.block_number)

    # views
    def version(self): return self.engine.version
    def now(self): return self.engine.now
    def block_number(self): return self.engine.block_number
    def eth_balance(self): return 10 ** 18

    def tasks(self, taskid):
        t = self.engine.tasks.get(taskid.lower())
        if t is None:
            return TaskView("0x" + "00" * 32, 0, ZERO_ADDR, 0, 0, "0x")
        return TaskView(t.model, t.fee, t.owner, t.blocktime, t.version, t.cid)

    def solutions(self, taskid):
        s = self.engine.solutions.get(taskid.lower())
        if s is None:
            return SolutionView(ZERO_ADDR, 0, False, "0x")
        return SolutionView(s.validator, s.blocktime, s.claimed, s.cid)

    def contestations(self, taskid):
        c = self.engine.contestations.get(taskid.lower())
        if c is None:
            return ContestationView(ZERO_ADDR, 0, 0, 0)
        return ContestationView(c.validator, c.blocktime, c.finish_start_index, c.slash_amount)

    def contestation_voted(self, taskid, addr):
        return bool(self.engine.contestation_voted.get((taskid.lower(), addr.lower())))

    def validator_staked(self, addr=None):
        return self.engine.validator(addr or self.address).staked

    def get_validator_minimum(self): return self.engine.get_validator_minimum()
    def token_balance(self, addr=None): return self.engine.token.balance_of(addr or self.address)
    def token_allowance(self, owner, spender): return self.engine.token.allowance(owner, spender)

    def task_input_bytes(self, taskid, txid):
        """The reference ABI-decodes submitTask calldata from getTransaction(txid) (index.ts:151-159)."""
        return self.engine.task_inputs[taskid.lower()]

    # transactions
    def token_approve(self, spender, amount):
        self.engine.token.approve(self.address, spender, amount)
        return Receipt(self.engine._txhash(), self.engine.block_number)

    def validator_deposit(self, validator, amount): return self._tx(self.engine.validator_deposit, self.address, validator, amount)

    def submit_task(self, version, owner, model, fee, input_):
        try:
            return self.engine.submit_task(self.address, version, owner, model, fee, input_)
        except Revert as e:
            raise ChainError(str(e)) from e

    def signal_commitment(self, commitment): return self._tx(self.engine.signal_commitment, self.address, commitment)
    def submit_solution(self, taskid, cid): return self._tx(self.engine.submit_solution, self.address, taskid, cid)
    def claim_solution(self, taskid): return self._tx(self.engine.claim_solution, self.address, taskid)
    def submit_contestation(self, taskid): return self._tx(self.engine.submit_contestation, self.address, taskid)
    def vote_on_contestation(self, taskid, yea): return self._tx(self.engine.vote_on_contestation, self.address, taskid, yea)
    def contestation_vote_finish(self, taskid, amnt): return self._tx(self.engine.contestation_vote_finish, self.address, taskid, amnt)

    def subscribe(self, handler): self.engine.subscribe(handler)
    def poll_events(self): return self.engine.deliver()
