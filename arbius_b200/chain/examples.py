"""Integration examples: the thirteen minimal contracts of the reference that call the Engine on behalf of a contract
address (reference: contract/contracts/Example/*.sol — SubmitTask, SubmitSolution, ClaimSolution, SubmitContestation,
VoteOnContestation, FinishContestationVote, RetractTask, RegisterModel, Lookup{TaskCID, SolutionCID, ModelAddress,
ValidatorStakedBalance, ContestationValidator}). Here an "example contract" is an address that owns tokens and calls
the engine state machine as ``msg.sender``; each class keeps the constructor arguments and the single method of its
Solidity counterpart, so a dapp author can see the call sequence (approve max, then call) and test against EngineSim.
"""
from __future__ import annotations

from .engine_sim import MAX_UINT256, UNIT, EngineSim


class _Integration:
    def __init__(self, arbius: EngineSim, address: str):
        self.arbius, self.address = arbius, address.lower()

    def _approve_max(self) -> None:
        self.arbius.token.approve(self.address, self.arbius.address, MAX_UINT256)


class SubmitTask(_Integration):
    """SubmitTask.sol:6-34 — approve max, submit with a fee of 0.1 AIUS, the contract itself as task owner"""

    def __init__(self, arbius, address, model: str, input_: bytes):
        super().__init__(arbius, address)
        self.model, self.input = model, input_

    def submit_task(self) -> str:
        self._approve_max()
        return self.arbius.submit_task(self.address, 0, self.address, self.model, UNIT // 10, self.input)


class SubmitSolution(_Integration):
    """SubmitSolution.sol:5-19 — commitment first, solution in a later block"""

    def signal_commitment(self, taskid: str, cid: str) -> None:
        self.arbius.signal_commitment(self.address, self.arbius.generate_commitment(self.address, taskid, cid))

    def submit_solution(self, taskid: str, cid: str) -> None:
        self.arbius.submit_solution(self.address, taskid, cid)


class ClaimSolution(_Integration):
    def claim_solution(self, taskid: str) -> None:
        self.arbius.claim_solution(self.address, taskid)


class SubmitContestation(_Integration):
    def submit_contestation(self, taskid: str) -> None:
        self.arbius.submit_contestation(self.address, taskid)


class VoteOnContestation(_Integration):
    def vote_on_contestation(self, taskid: str, yea: bool) -> None:
        self.arbius.vote_on_contestation(self.address, taskid, yea)


class FinishContestationVote(_Integration):
    def finish(self, taskid: str, amnt: int) -> None:
        self.arbius.contestation_vote_finish(self.address, taskid, amnt)


class RetractTask(_Integration):
    def retract_task(self, taskid: str) -> None:
        self.arbius.retract_task(self.address, taskid)


class RegisterModel(_Integration):
    def register_model(self, addr: str, fee: int, template: bytes) -> str:
        return self.arbius.register_model(self.address, addr, fee, template)


class Lookup(_Integration):
    """the five Lookup*.sol view helpers"""

    def task_cid(self, taskid: str) -> str:
        return self.arbius.tasks[taskid].cid

    def solution_cid(self, taskid: str) -> str:
        return self.arbius.solutions[taskid].cid

    def model_address(self, model: str) -> str:
        return self.arbius.models[model].addr

    def validator_staked_balance(self, validator: str) -> int:
        return self.arbius.validator(validator).staked

    def contestation_validator(self, taskid: str) -> str:
        return self.arbius.contestations[taskid].validator
