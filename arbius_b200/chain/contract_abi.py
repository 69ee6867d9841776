"""Function and event tables of the deployed contracts, as (name -> input types, output types) and
(name -> [(arg, type, indexed)]). Same surface as the Hardhat artifacts the reference ships
(reference: miner/src/artifacts/contracts/EngineV1.sol/EngineV1.json — 77 functions, 31 events; BaseTokenV1.json),
restated as plain tuples so selectors/topics are computed here instead of being read from a JSON artifact.
"""
from __future__ import annotations

from typing import Any, Dict, List, Sequence, Tuple

from ..protocol import abi
from ..protocol.keccak import keccak256

MODEL_T = "(uint256,address,uint256,bytes)"
TASK_T = "(bytes32,uint256,address,uint64,uint8,bytes)"

# name: (inputs, outputs, mutates)
ENGINE_FUNCTIONS: Dict[str, Tuple[List[str], List[str], bool]] = {
    # views over storage
    "accruedFees": ([], ["uint256"], False), "baseToken": ([], ["address"], False),
    "commitments": (["bytes32"], ["uint256"], False),
    "contestationVoteNays": (["bytes32", "uint256"], ["address"], False),
    "contestationVoteYeas": (["bytes32", "uint256"], ["address"], False),
    "contestationVoted": (["bytes32", "address"], ["bool"], False),
    "contestationVotedIndex": (["bytes32"], ["uint256"], False),
    "contestations": (["bytes32"], ["address", "uint64", "uint32", "uint256"], False),
    "exitValidatorMinUnlockTime": ([], ["uint256"], False),
    "maxContestationValidatorStakeSince": ([], ["uint256"], False),
    "minClaimSolutionTime": ([], ["uint256"], False), "minContestationVotePeriodTime": ([], ["uint256"], False),
    "minRetractionWaitTime": ([], ["uint256"], False),
    "models": (["bytes32"], ["uint256", "address", "uint256", "bytes"], False),
    "owner": ([], ["address"], False), "paused": ([], ["bool"], False), "pauser": ([], ["address"], False),
    "pendingValidatorWithdrawRequests": (["address", "uint256"], ["uint256", "uint256"], False),
    "pendingValidatorWithdrawRequestsCount": (["address"], ["uint256"], False),
    "prevhash": ([], ["bytes32"], False), "retractionFeePercentage": ([], ["uint256"], False),
    "slashAmountPercentage": ([], ["uint256"], False), "solutionFeePercentage": ([], ["uint256"], False),
    "solutions": (["bytes32"], ["address", "uint64", "bool", "bytes"], False),
    "startBlockTime": ([], ["uint64"], False),
    "tasks": (["bytes32"], ["bytes32", "uint256", "address", "uint64", "uint8", "bytes"], False),
    "treasury": ([], ["address"], False), "treasuryRewardPercentage": ([], ["uint256"], False),
    "validatorMinimumPercentage": ([], ["uint256"], False),
    "validatorWithdrawPendingAmount": (["address"], ["uint256"], False),
    "validators": (["address"], ["uint256", "uint256", "address"], False), "version": ([], ["uint256"], False),
    # computed views / pure helpers
    "diffMul": (["uint256", "uint256"], ["uint256"], False),
    "generateCommitment": (["address", "bytes32", "bytes"], ["bytes32"], False),
    "generateIPFSCID": (["bytes"], ["bytes"], False), "getPsuedoTotalSupply": ([], ["uint256"], False),
    "getReward": ([], ["uint256"], False), "getSlashAmount": ([], ["uint256"], False),
    "getValidatorMinimum": ([], ["uint256"], False),
    "hashModel": ([MODEL_T, "address"], ["bytes32"], False),
    "hashTask": ([TASK_T, "address", "bytes32"], ["bytes32"], False),
    "reward": (["uint256", "uint256"], ["uint256"], False), "targetTs": (["uint256"], ["uint256"], False),
    "validatorCanVote": (["address", "bytes32"], ["uint256"], False),
    # transactions
    "cancelValidatorWithdraw": (["uint256"], [], True), "claimSolution": (["bytes32"], [], True),
    "contestationVoteFinish": (["bytes32", "uint32"], [], True), "initialize": (["address", "address"], [], True),
    "initiateValidatorWithdraw": (["uint256"], ["uint256"], True),
    "registerModel": (["address", "uint256", "bytes"], ["bytes32"], True), "renounceOwnership": ([], [], True),
    "retractTask": (["bytes32"], [], True), "setExitValidatorMinUnlockTime": (["uint256"], [], True),
    "setMaxContestationValidatorStakeSince": (["uint256"], [], True), "setMinClaimSolutionTime": (["uint256"], [], True),
    "setMinContestationVotePeriodTime": (["uint256"], [], True), "setMinRetractionWaitTime": (["uint256"], [], True),
    "setPaused": (["bool"], [], True), "setRetractionFeePercentage": (["uint256"], [], True),
    "setSlashAmountPercentage": (["uint256"], [], True), "setSolutionFeePercentage": (["uint256"], [], True),
    "setSolutionMineableRate": (["bytes32", "uint256"], [], True), "setTreasuryRewardPercentage": (["uint256"], [], True),
    "setValidatorMinimumPercentage": (["uint256"], [], True), "setVersion": (["uint256"], [], True),
    "signalCommitment": (["bytes32"], [], True), "signalSupport": (["bytes32", "bool"], [], True),
    "submitContestation": (["bytes32"], [], True), "submitSolution": (["bytes32", "bytes"], [], True),
    "submitTask": (["uint8", "address", "bytes32", "uint256", "bytes"], ["bytes32"], True),
    "transferOwnership": (["address"], [], True), "transferPauser": (["address"], [], True),
    "transferTreasury": (["address"], [], True), "validatorDeposit": (["address", "uint256"], [], True),
    "validatorWithdraw": (["uint256", "address"], [], True), "voteOnContestation": (["bytes32", "bool"], [], True),
    "withdrawAccruedFees": ([], [], True),
}

_AMT = [("amount", "uint256", True)]
ENGINE_EVENTS: Dict[str, List[Tuple[str, str, bool]]] = {
    "ModelRegistered": [("id", "bytes32", True)],
    "ValidatorDeposit": [("addr", "address", True), ("validator", "address", True), ("amount", "uint256", False)],
    "ValidatorWithdrawInitiated": [("addr", "address", True), ("count", "uint256", True), ("unlockTime", "uint256", False),
                                   ("amount", "uint256", False)],
    "ValidatorWithdrawCancelled": [("addr", "address", True), ("count", "uint256", True)],
    "ValidatorWithdraw": [("addr", "address", True), ("to", "address", True), ("count", "uint256", True),
                          ("amount", "uint256", False)],
    "TaskSubmitted": [("id", "bytes32", True), ("model", "bytes32", True), ("fee", "uint256", False),
                      ("sender", "address", True)],
    "TaskRetracted": [("id", "bytes32", True)],
    "SignalCommitment": [("addr", "address", True), ("commitment", "bytes32", True)],
    "SolutionSubmitted": [("addr", "address", True), ("task", "bytes32", True)],
    "SolutionClaimed": [("addr", "address", True), ("task", "bytes32", True)],
    "ContestationSubmitted": [("addr", "address", True), ("task", "bytes32", True)],
    "ContestationVote": [("addr", "address", True), ("task", "bytes32", True), ("yea", "bool", False)],
    "ContestationVoteFinish": [("id", "bytes32", True), ("start_idx", "uint32", True), ("end_idx", "uint32", False)],
    "SignalSupport": [("addr", "address", True), ("model", "bytes32", True), ("supported", "bool", False)],
    "TreasuryTransferred": [("to", "address", True)], "PauserTransferred": [("to", "address", True)],
    "PausedChanged": [("paused", "bool", True)],
    "SolutionMineableRateChange": [("id", "bytes32", True), ("rate", "uint256", False)],
    "VersionChanged": [("version", "uint256", False)],
    "ValidatorMinimumPercentageChanged": _AMT, "SlashAmountPercentageChanged": _AMT, "SolutionFeePercentageChanged": _AMT,
    "RetractionFeePercentageChanged": _AMT, "TreasuryRewardPercentageChanged": _AMT, "MinClaimSolutionTimeChanged": _AMT,
    "MinRetractionWaitTimeChanged": _AMT, "MinContestationVotePeriodTimeChanged": _AMT,
    "MaxContestationValidatorStakeSinceChanged": _AMT, "ExitValidatorMinUnlockTimeChanged": _AMT,
    "OwnershipTransferred": [("previousOwner", "address", True), ("newOwner", "address", True)],
    "Initialized": [("version", "uint8", False)],
}

TOKEN_FUNCTIONS: Dict[str, Tuple[List[str], List[str], bool]] = {
    "name": ([], ["string"], False), "symbol": ([], ["string"], False), "decimals": ([], ["uint8"], False),
    "totalSupply": ([], ["uint256"], False), "balanceOf": (["address"], ["uint256"], False),
    "allowance": (["address", "address"], ["uint256"], False), "l2Gateway": ([], ["address"], False),
    "l1Address": ([], ["address"], False), "owner": ([], ["address"], False),
    "delegates": (["address"], ["address"], False), "getVotes": (["address"], ["uint256"], False),
    "getPastVotes": (["address", "uint256"], ["uint256"], False),
    "approve": (["address", "uint256"], ["bool"], True), "transfer": (["address", "uint256"], ["bool"], True),
    "transferFrom": (["address", "address", "uint256"], ["bool"], True), "delegate": (["address"], [], True),
    "bridgeMint": (["address", "uint256"], [], True), "bridgeBurn": (["address", "uint256"], [], True),
    "transferOwnership": (["address"], [], True),
}
TOKEN_EVENTS: Dict[str, List[Tuple[str, str, bool]]] = {
    "Transfer": [("from", "address", True), ("to", "address", True), ("value", "uint256", False)],
    "Approval": [("owner", "address", True), ("spender", "address", True), ("value", "uint256", False)],
    "DelegateChanged": [("delegator", "address", True), ("fromDelegate", "address", True), ("toDelegate", "address", True)],
    "DelegateVotesChanged": [("delegate", "address", True), ("previousBalance", "uint256", False),
                             ("newBalance", "uint256", False)],
}

GOVERNOR_FUNCTIONS: Dict[str, Tuple[List[str], List[str], bool]] = {
    "name": ([], ["string"], False), "votingDelay": ([], ["uint256"], False), "votingPeriod": ([], ["uint256"], False),
    "proposalThreshold": ([], ["uint256"], False), "quorum": (["uint256"], ["uint256"], False),
    "state": (["uint256"], ["uint8"], False), "proposalSnapshot": (["uint256"], ["uint256"], False),
    "proposalDeadline": (["uint256"], ["uint256"], False), "proposalEta": (["uint256"], ["uint256"], False),
    "proposalVotes": (["uint256"], ["uint256", "uint256", "uint256"], False),
    "hasVoted": (["uint256", "address"], ["bool"], False), "proposalsCreated": (["uint256"], ["uint256"], False),
    "proposalsCreatedLength": ([], ["uint256"], False),
    "descriptionHashes": (["uint256"], ["bytes32"], False), "descriptionCids": (["uint256"], ["bytes"], False),
    "hashProposal": (["address[]", "uint256[]", "bytes[]", "bytes32"], ["uint256"], False),
    "timelock": ([], ["address"], False),
    "getActions": (["uint256"], ["address[]", "uint256[]", "string[]", "bytes[]"], False),
    "proposals": (["uint256"], ["uint256", "address", "uint256", "uint256", "uint256", "uint256", "uint256", "uint256",
                                "bool", "bool"], False),
    "propose": (["address[]", "uint256[]", "bytes[]", "string"], ["uint256"], True),
    "castVote": (["uint256", "uint8"], ["uint256"], True),
    "castVoteWithReason": (["uint256", "uint8", "string"], ["uint256"], True),
    "queue": (["address[]", "uint256[]", "bytes[]", "bytes32"], ["uint256"], True),
    "execute": (["address[]", "uint256[]", "bytes[]", "bytes32"], ["uint256"], True),
    "cancel": (["address[]", "uint256[]", "bytes[]", "bytes32"], ["uint256"], True),
}
GOVERNOR_EVENTS: Dict[str, List[Tuple[str, str, bool]]] = {
    "ProposalCreated": [("proposalId", "uint256", False), ("proposer", "address", False), ("targets", "address[]", False),
                        ("values", "uint256[]", False), ("signatures", "string[]", False), ("calldatas", "bytes[]", False),
                        ("startBlock", "uint256", False), ("endBlock", "uint256", False), ("description", "string", False)],
    "VoteCast": [("voter", "address", True), ("proposalId", "uint256", False), ("support", "uint8", False),
                 ("weight", "uint256", False), ("reason", "string", False)],
    "ProposalQueued": [("proposalId", "uint256", False), ("eta", "uint256", False)],
    "ProposalExecuted": [("proposalId", "uint256", False)], "ProposalCanceled": [("proposalId", "uint256", False)],
}


def signature(name: str, inputs: Sequence[str]) -> str:
    return f"{name}({','.join(inputs)})"


def selector_of(name: str, table: Dict[str, tuple]) -> bytes:
    return keccak256(signature(name, table[name][0]).encode())[:4]


def selector_table(table: Dict[str, tuple]) -> Dict[bytes, str]:
    return {selector_of(n, table): n for n in table}


def encode_call(table: Dict[str, tuple], name: str, args: Sequence[Any]) -> bytes:
    return selector_of(name, table) + abi.encode(table[name][0], list(args))


def decode_result(table: Dict[str, tuple], name: str, data: bytes) -> list:
    return abi.decode(table[name][1], data)


def event_topic(name: str, table: Dict[str, list]) -> bytes:
    return keccak256(signature(name, [t for _, t, _ in table[name]]).encode())


def _topic_word(t: str, v: Any) -> bytes:
    if t in ("bytes", "string") or t.endswith("]"):
        raw = v.encode() if isinstance(v, str) and t == "string" else abi.encode([t], [v])
        return keccak256(raw)
    return abi.encode([t], [v])


def encode_log(table: Dict[str, list], name: str, args: Dict[str, Any]) -> Tuple[List[str], str]:
    """(topics, data) hex strings of one emitted event"""
    spec = table[name]
    topics = [event_topic(name, table)] + [_topic_word(t, args[a]) for a, t, idx in spec if idx]
    plain = [(t, args[a]) for a, t, idx in spec if not idx]
    data = abi.encode([t for t, _ in plain], [v for _, v in plain])
    return ["0x" + t.hex() for t in topics], "0x" + data.hex()


def decode_log(table: Dict[str, list], topics: Sequence[str], data: str):
    """-> (event name, args dict) or None if topic0 is not in the table"""
    by_topic = {"0x" + event_topic(n, table).hex(): n for n in table}
    name = by_topic.get(topics[0].lower()) if topics else None
    if name is None:
        return None
    spec = table[name]
    out: Dict[str, Any] = {}
    ti = 1
    for a, t, idx in spec:
        if idx:
            word = bytes.fromhex(topics[ti][2:])
            ti += 1
            out[a] = word if (t in ("bytes", "string") or t.endswith("]")) else abi.decode([t], word)[0]
    plain = [(a, t) for a, t, idx in spec if not idx]
    vals = abi.decode([t for _, t in plain], bytes.fromhex(data[2:])) if plain else []
    for (a, _), v in zip(plain, vals):
        out[a] = v
    for a, t, _ in spec:
        if t.startswith("bytes") and t != "bytes" and isinstance(out.get(a), (bytes, bytearray)):
            out[a] = "0x" + bytes(out[a]).hex()
    return name, out
