"""DevChain — the protocol contracts behind an Ethereum execution-client surface, in process.

This is the stand-in for ``npx hardhat node`` + the deploy scripts (reference: setup_local.sh:1-23,
contract/scripts/003-deploy-core-basic.ts:7-112): signed raw transactions in, ABI-encoded return data, receipts and
logs out, with Hardhat's automine semantics (one block per transaction) and its time-travel RPCs. Contract code is not
EVM bytecode but the Python state machines of engine_sim.py / governance.py, addressed by contract address + selector.
``chain/rpc_server.py`` exposes it over HTTP JSON-RPC; ``JsonRpcChainClient`` and the CLI talk to either.
"""
from __future__ import annotations

import copy
import re
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple

from ..protocol import abi
from ..protocol.abi import hash_model, hash_task
from ..protocol.keccak import keccak256
from . import contract_abi as CA
from . import prbmath
from .engine_sim import STARTING_ENGINE_TOKEN_AMOUNT, ZERO_ADDR, EngineSim, Revert, _addr, _hex
from .eth import Transaction, decode_raw_transaction
from .governance import GovernorSim, TimelockSim, VotesToken

_Z32 = "0x" + "00" * 32


class RpcError(Exception):
    def __init__(self, code: int, message: str, data: Optional[str] = None):
        super().__init__(message)
        self.code, self.message, self.data = code, message, data


def revert_data(reason: str) -> str:
    return "0x08c379a0" + abi.encode(["string"], [reason]).hex()


def _snake(name: str) -> str:
    return re.sub(r"(?<!^)(?=[A-Z])", "_", name).lower()


@dataclass
class Block:
    number: int
    timestamp: int
    tx_hashes: List[str] = field(default_factory=list)

    @property
    def hash(self) -> str:
        return "0x" + keccak256(b"devchain-block" + self.number.to_bytes(8, "big")).hex()


class DevChain:
    def __init__(self, chain_id: int = 31337, deployer: str = "0x" + "0a" * 20, funded: Optional[Dict[str, int]] = None,
                 timelock_delay: int = 0, start_time: int = 1_700_000_000,
                 engine_balance: int = STARTING_ENGINE_TOKEN_AMOUNT - 1):
        self.chain_id = chain_id
        token = VotesToken(owner=deployer)
        self.engine = EngineSim(token=token, owner=deployer, start_time=start_time, chain_id=chain_id)
        token.clock = lambda: self.engine.block_number
        self.token = token
        token.mint(self.engine.address, engine_balance)
        for a, amt in (funded or {}).items():
            token.mint(a, amt)
        self.timelock = TimelockSim(timelock_delay, lambda: self.engine.now)
        self.governor = GovernorSim(token, self.timelock, lambda: self.engine.block_number, self._internal_call)
        self.timelock.grant_proposer(self.governor.address)
        self.nonces: Dict[str, int] = {}
        self.blocks: List[Block] = [Block(self.engine.block_number, self.engine.now)]
        self.txs: Dict[str, dict] = {}
        self.receipts: Dict[str, dict] = {}
        self.logs: List[dict] = []
        self._cursors = {"engine": 0, "token": 0, "governor": 0}
        self._snapshots: List[Any] = []
        self._sel = {self.engine.address: (CA.selector_table(CA.ENGINE_FUNCTIONS), CA.ENGINE_FUNCTIONS, self._engine_call),
                     self.token.address: (CA.selector_table(CA.TOKEN_FUNCTIONS), CA.TOKEN_FUNCTIONS, self._token_call),
                     self.governor.address: (CA.selector_table(CA.GOVERNOR_FUNCTIONS), CA.GOVERNOR_FUNCTIONS,
                                             self._governor_call)}

    # ------------------------------------------------------------------------------------ addresses
    @property
    def addresses(self) -> dict:
        return {"engineAddress": self.engine.address, "baseTokenAddress": self.token.address,
                "governorAddress": self.governor.address, "timelockAddress": self.timelock.address}

    # ------------------------------------------------------------------------------------ state copy
    def _state(self):
        e = self.engine
        listeners, e._listeners = e._listeners, []
        clock, self.token.clock = self.token.clock, None
        g_call, g_block, self.governor.call, self.governor.block = self.governor.call, self.governor.block, None, None
        t_now, self.timelock.now = self.timelock.now, None
        g_tok, g_tl, self.governor.token, self.governor.timelock = self.governor.token, self.governor.timelock, None, None
        try:
            snap = copy.deepcopy((e.__dict__, self.governor.__dict__, self.timelock.__dict__, self.nonces))
        finally:
            e._listeners, self.token.clock = listeners, clock
            self.governor.call, self.governor.block, self.timelock.now = g_call, g_block, t_now
            self.governor.token, self.governor.timelock = g_tok, g_tl
        return snap

    def _restore(self, snap) -> None:
        e = self.engine
        listeners = e._listeners
        ed, gd, td, nonces = copy.deepcopy(snap)
        e.__dict__.update(ed)
        e._listeners = listeners
        self.token = e.token
        self.token.clock = lambda: self.engine.block_number
        for k, v in gd.items():
            if k not in ("call", "block", "token", "timelock"):
                setattr(self.governor, k, v)
        for k, v in td.items():
            if k != "now":
                setattr(self.timelock, k, v)
        self.governor.token = self.token
        self.nonces = nonces

    # ------------------------------------------------------------------------------------ dispatch
    def _decode(self, to: str, data: bytes):
        ent = self._sel.get(_addr(to))
        if ent is None:
            raise Revert("call to a non-contract account")
        table, funcs, handler = ent
        name = table.get(bytes(data[:4]))
        if name is None:
            raise Revert("function selector was not recognized")
        return name, abi.decode(funcs[name][0], bytes(data[4:])), funcs[name], handler

    def _internal_call(self, sender: str, to: str, value: int, data: bytes) -> None:
        name, args, _, handler = self._decode(to, data)
        handler(sender, name, args)

    def call(self, sender: str, to: str, data: bytes) -> bytes:
        name, args, spec, handler = self._decode(to, data)
        out = handler(_addr(sender), name, args)
        if not spec[1]:
            return b""
        vals = list(out) if isinstance(out, (tuple, list)) and len(spec[1]) > 1 else [out]
        return abi.encode(spec[1], vals)

    # --- engine
    _ENGINE_ATTRS = {"accruedFees": "accrued_fees", "exitValidatorMinUnlockTime": "exit_validator_min_unlock_time",
                     "maxContestationValidatorStakeSince": "max_contestation_validator_stake_since",
                     "minClaimSolutionTime": "min_claim_solution_time",
                     "minContestationVotePeriodTime": "min_contestation_vote_period_time",
                     "minRetractionWaitTime": "min_retraction_wait_time", "owner": "owner", "paused": "paused",
                     "pauser": "pauser", "prevhash": "prevhash", "retractionFeePercentage": "retraction_fee_percentage",
                     "slashAmountPercentage": "slash_amount_percentage", "solutionFeePercentage": "solution_fee_percentage",
                     "startBlockTime": "start_block_time", "treasury": "treasury",
                     "treasuryRewardPercentage": "treasury_reward_percentage",
                     "validatorMinimumPercentage": "validator_minimum_percentage", "version": "version"}

    def _engine_call(self, sender: str, name: str, a: list):
        e = self.engine
        if name in self._ENGINE_ATTRS:
            return getattr(e, self._ENGINE_ATTRS[name])
        h = (lambda i: _hex(a[i]))
        if name == "baseToken":
            return e.token.address
        if name == "commitments":
            return e.commitments.get(h(0), 0)
        if name in ("contestationVoteNays", "contestationVoteYeas"):
            arr = (e.contestation_nays if name.endswith("Nays") else e.contestation_yeas).get(h(0), [])
            if a[1] >= len(arr):
                raise Revert("")
            return arr[a[1]]
        if name == "contestationVoted":
            return bool(e.contestation_voted.get((h(0), _addr(a[1]))))
        if name == "contestationVotedIndex":
            return 0
        if name == "contestations":
            c = e.contestations.get(h(0))
            return (c.validator, c.blocktime, c.finish_start_index, c.slash_amount) if c else (ZERO_ADDR, 0, 0, 0)
        if name == "models":
            m = e.models.get(h(0))
            return (m.fee, m.addr, m.rate, m.cid) if m else (0, ZERO_ADDR, 0, b"")
        if name == "solutions":
            s = e.solutions.get(h(0))
            return (s.validator, s.blocktime, s.claimed, s.cid) if s else (ZERO_ADDR, 0, False, b"")
        if name == "tasks":
            t = e.tasks.get(h(0))
            return (t.model, t.fee, t.owner, t.blocktime, t.version, t.cid) if t else (_Z32, 0, ZERO_ADDR, 0, 0, b"")
        if name == "validators":
            v = e.validator(a[0])
            return v.staked, v.since, v.addr
        if name == "pendingValidatorWithdrawRequests":
            return e.withdraw_requests.get((_addr(a[0]), a[1]), (0, 0))
        if name == "pendingValidatorWithdrawRequestsCount":
            return e.withdraw_counts.get(_addr(a[0]), 0)
        if name == "validatorWithdrawPendingAmount":
            return e.withdraw_pending.get(_addr(a[0]), 0)
        if name == "diffMul":
            return prbmath.diff_mul(a[0], a[1])
        if name == "reward":
            return prbmath.reward(a[0], a[1])
        if name == "targetTs":
            return prbmath.target_ts(a[0])
        if name == "generateCommitment":
            return e.generate_commitment(a[0], h(1), _hex(a[2]))
        if name == "generateIPFSCID":
            return e.generate_ipfs_cid(a[0])
        if name == "hashModel":
            fee, addr, rate, cid = a[0]
            return hash_model(addr, fee, rate, cid, a[1])
        if name == "hashTask":
            model, fee, _owner, _bt, _ver, cid = a[0]
            return hash_task(_hex(model), fee, cid, a[1], h(2))
        if name in ("getPsuedoTotalSupply", "getReward", "getSlashAmount", "getValidatorMinimum"):
            return getattr(e, _snake(name))()
        if name == "validatorCanVote":
            return e.validator_can_vote(a[0], h(1))
        if name == "initialize":
            raise Revert("Initializable: contract is already initialized")
        if name == "renounceOwnership":
            e._only_owner(sender)
            e.owner = ZERO_ADDR
            return None
        args = [(_hex(v) if isinstance(v, (bytes, bytearray)) and t == "bytes32" else v)
                for v, t in zip(a, CA.ENGINE_FUNCTIONS[name][0])]
        if name == "submitSolution":
            args[1] = _hex(args[1])
        return getattr(e, _snake(name))(sender, *args)

    # --- token
    def _token_call(self, sender: str, name: str, a: list):
        t = self.token
        simple = {"name": t.name, "symbol": t.symbol, "decimals": t.decimals, "totalSupply": t.total_supply,
                  "l2Gateway": t.l2_gateway, "l1Address": t.l1_address, "owner": t.owner}
        if name in simple:
            return simple[name]
        if name == "balanceOf":
            return t.balance_of(a[0])
        if name == "allowance":
            return t.allowance(a[0], a[1])
        if name == "delegates":
            return t.delegates(a[0])
        if name == "getVotes":
            return t.get_votes(a[0])
        if name == "getPastVotes":
            return t.get_past_votes(a[0], a[1])
        if name == "approve":
            t.approve(sender, a[0], a[1])
            t.events.append(("Approval", {"owner": sender, "spender": _addr(a[0]), "value": a[1]}))
            return True
        if name == "transfer":
            t.transfer(sender, a[0], a[1])
            t.events.append(("Transfer", {"from": sender, "to": _addr(a[0]), "value": a[1]}))
            return True
        if name == "transferFrom":
            t.transfer_from(sender, a[0], a[1], a[2])
            t.events.append(("Transfer", {"from": _addr(a[0]), "to": _addr(a[1]), "value": a[2]}))
            return True
        if name == "delegate":
            return t.delegate(sender, a[0])
        if name == "bridgeMint":
            return t.bridge_mint(sender, a[0], a[1])
        if name == "bridgeBurn":
            return t.bridge_burn(sender, a[0], a[1])
        if name == "transferOwnership":
            return t.transfer_ownership(sender, a[0])
        raise Revert("function selector was not recognized")

    # --- governor
    def _governor_call(self, sender: str, name: str, a: list):
        g = self.governor
        if name == "name":
            return g.name
        if name in ("votingDelay", "votingPeriod", "proposalThreshold"):
            return getattr(g, _snake(name))
        if name == "quorum":
            return g.quorum(a[0])
        if name == "state":
            return g.state(a[0])
        if name in ("proposalSnapshot", "proposalDeadline"):
            p = g.proposals.get(a[0])
            if p is None:
                raise Revert("Governor: unknown proposal id")
            return p.vote_start if name == "proposalSnapshot" else p.vote_end
        if name == "proposalEta":
            return g.proposal_eta(a[0])
        if name == "proposalVotes":
            p = g.proposals.get(a[0])
            return tuple(p.votes) if p else (0, 0, 0)
        if name == "hasVoted":
            p = g.proposals.get(a[0])
            return bool(p and p.voted.get(_addr(a[1])))
        if name == "proposalsCreated":
            if a[0] >= len(g.proposals_created):
                raise Revert("")
            return g.proposals_created[a[0]]
        if name == "proposalsCreatedLength":
            return len(g.proposals_created)
        if name == "descriptionHashes":
            return g.description_hashes.get(a[0], b"\x00" * 32)
        if name == "descriptionCids":
            return g.description_cids.get(a[0], b"")
        if name == "hashProposal":
            return g.hash_proposal(a[0], a[1], a[2], a[3])
        if name == "timelock":
            return self.timelock.address
        if name in ("getActions", "proposals"):
            p = g.proposals.get(a[0])
            if p is None:
                raise Revert("Governor: unknown proposal id")
            if name == "getActions":
                return p.targets, p.values, [""] * len(p.targets), p.calldatas
            return (a[0], p.proposer, g.proposal_eta(a[0]), p.vote_start, p.vote_end, p.votes[1], p.votes[0], p.votes[2],
                    g.state(a[0]) == 2, g.state(a[0]) == 7)
        if name == "propose":
            return g.propose(sender, a[0], a[1], a[2], a[3])
        if name == "castVote":
            return g.cast_vote(sender, a[0], a[1])
        if name == "castVoteWithReason":
            return g.cast_vote(sender, a[0], a[1], a[2])
        if name in ("queue", "execute", "cancel"):
            return getattr(g, name)(sender, a[0], a[1], a[2], a[3])
        raise Revert("function selector was not recognized")

    # ------------------------------------------------------------------------------------ blocks / txs
    def _new_block(self, tx_hash: Optional[str] = None) -> Block:
        e = self.engine
        if e.block_number <= self.blocks[-1].number:
            e.block_number = self.blocks[-1].number + 1
        if e.now <= self.blocks[-1].timestamp:
            e.now = self.blocks[-1].timestamp + 1
        b = Block(e.block_number, e.now, [tx_hash] if tx_hash else [])
        self.blocks.append(b)
        return b

    def _collect_logs(self, tx_hash: str, block: Block) -> List[dict]:
        out = []
        sources = (("engine", self.engine.events, self.engine.address, CA.ENGINE_EVENTS,
                    lambda ev: (ev.name, ev.args)),
                   ("token", self.token.events, self.token.address, CA.TOKEN_EVENTS, lambda ev: ev),
                   ("governor", self.governor.events, self.governor.address, CA.GOVERNOR_EVENTS, lambda ev: ev))
        for key, events, address, table, unpack in sources:
            while self._cursors[key] < len(events):
                name, args = unpack(events[self._cursors[key]])
                self._cursors[key] += 1
                if name not in table:
                    continue
                topics, data = CA.encode_log(table, name, args)
                out.append({"address": address, "topics": topics, "data": data, "blockNumber": hex(block.number),
                            "blockHash": block.hash, "transactionHash": tx_hash, "transactionIndex": "0x0",
                            "logIndex": hex(len(out)), "removed": False})
        return out

    def _trim_events(self) -> None:
        """after a revert the event lists were restored; cursors may point past their ends"""
        self._cursors["engine"] = min(self._cursors["engine"], len(self.engine.events))
        self._cursors["token"] = min(self._cursors["token"], len(self.token.events))
        self._cursors["governor"] = min(self._cursors["governor"], len(self.governor.events))

    def send_raw_transaction(self, raw: bytes) -> str:
        try:
            tx = decode_raw_transaction(raw)
        except (ValueError, IndexError) as err:
            raise RpcError(-32000, f"invalid transaction: {err}") from err
        if tx.chain_id != self.chain_id:
            raise RpcError(-32000, f"invalid chain id: have {tx.chain_id}, want {self.chain_id}")
        expected = self.nonces.get(tx.sender, 0)
        if tx.nonce != expected:
            raise RpcError(-32000, f"nonce too {'low' if tx.nonce < expected else 'high'}: expected {expected}, got {tx.nonce}")
        if tx.to is None:
            raise RpcError(-32000, "contract creation is not supported: contracts are pre-deployed (see arbius_getAddresses)")
        snap = self._state()
        status, reason, ret = 1, None, b""
        try:
            ret = self.call(tx.sender, tx.to, tx.data) if tx.data else b""
        except Revert as err:
            status, reason = 0, str(err)
            self._restore(snap)
            self._trim_events()
        except (ValueError, IndexError, KeyError, TypeError) as err:       # malformed calldata
            status, reason = 0, f"invalid call: {err}"
            self._restore(snap)
            self._trim_events()
        self.nonces[tx.sender] = expected + 1
        block = self._new_block(tx.hash)
        logs = self._collect_logs(tx.hash, block) if status else []
        self.logs.extend(logs)
        self.txs[tx.hash] = {"hash": tx.hash, "from": tx.sender, "to": tx.to, "nonce": hex(tx.nonce), "value": hex(tx.value),
                             "input": "0x" + tx.data.hex(), "gas": hex(tx.gas), "blockNumber": hex(block.number),
                             "blockHash": block.hash, "transactionIndex": "0x0", "chainId": hex(self.chain_id),
                             "gasPrice": hex(tx.gas_price or tx.max_fee_per_gas or 0), "type": "0x2" if tx.is_1559 else "0x0"}
        self.receipts[tx.hash] = {"transactionHash": tx.hash, "transactionIndex": "0x0", "blockNumber": hex(block.number),
                                  "blockHash": block.hash, "from": tx.sender, "to": tx.to, "status": hex(status),
                                  "gasUsed": hex(21000 + 16 * len(tx.data)), "cumulativeGasUsed": hex(21000 + 16 * len(tx.data)),
                                  "effectiveGasPrice": hex(tx.gas_price or tx.max_fee_per_gas or 0), "contractAddress": None,
                                  "logs": logs, "logsBloom": "0x" + "00" * 256, "type": "0x2" if tx.is_1559 else "0x0",
                                  "revertReason": reason, "returnData": "0x" + ret.hex()}
        self.engine.deliver()
        return tx.hash

    def eth_call(self, sender: str, to: str, data: bytes) -> bytes:
        """read-only execution: state is rolled back afterwards, reverts surface as RpcError(3)"""
        snap = self._state()
        try:
            return self.call(sender or ZERO_ADDR, to, data)
        except Revert as err:
            raise RpcError(3, f"execution reverted: {err}", revert_data(str(err))) from err
        finally:
            self._restore(snap)
            self._trim_events()

    def get_logs(self, flt: dict) -> List[dict]:
        def num(v, default):
            if v in (None, "latest", "pending"):
                return default
            if v == "earliest":
                return 0
            return int(v, 16)
        lo, hi = num(flt.get("fromBlock"), self.blocks[-1].number), num(flt.get("toBlock"), self.blocks[-1].number)
        addr = flt.get("address")
        addrs = None if addr is None else {_addr(x) for x in ([addr] if isinstance(addr, str) else addr)}
        topics = flt.get("topics") or []
        out = []
        for lg in self.logs:
            bn = int(lg["blockNumber"], 16)
            if bn < lo or bn > hi or (addrs is not None and lg["address"] not in addrs):
                continue
            ok = True
            for i, want in enumerate(topics):
                if want is None:
                    continue
                opts = [want] if isinstance(want, str) else want
                if i >= len(lg["topics"]) or lg["topics"][i].lower() not in [o.lower() for o in opts]:
                    ok = False
                    break
            if ok:
                out.append(lg)
        return out

    # ------------------------------------------------------------------------------------ hardhat-style controls
    def mine(self, blocks: int = 1, interval: int = 1) -> None:
        for _ in range(blocks):
            self.engine.block_number += 1
            self.engine.now += interval
            self.blocks.append(Block(self.engine.block_number, self.engine.now))

    def increase_time(self, seconds: int) -> None:
        self.engine.now += seconds

    def snapshot(self) -> int:
        self._snapshots.append((self._state(), len(self.blocks), len(self.logs), dict(self._cursors),
                                set(self.txs), set(self.receipts)))
        return len(self._snapshots)

    def revert(self, snap_id: int) -> bool:
        if not 0 < snap_id <= len(self._snapshots):
            return False
        state, nblocks, nlogs, cursors, txs, receipts = self._snapshots[snap_id - 1]
        del self._snapshots[snap_id - 1:]
        self._restore(state)
        del self.blocks[nblocks:]
        del self.logs[nlogs:]
        self._cursors = cursors
        self.txs = {k: v for k, v in self.txs.items() if k in txs}
        self.receipts = {k: v for k, v in self.receipts.items() if k in receipts}
        return True

    def block_json(self, b: Block, full: bool = False) -> dict:
        return {"number": hex(b.number), "hash": b.hash, "timestamp": hex(b.timestamp), "parentHash": _Z32,
                "transactions": [self.txs[h] for h in b.tx_hashes] if full else list(b.tx_hashes),
                "gasLimit": hex(30_000_000), "gasUsed": "0x0", "baseFeePerGas": "0x0", "miner": ZERO_ADDR,
                "difficulty": "0x0", "extraData": "0x", "nonce": "0x0000000000000000", "size": "0x0"}

    def block_by_number(self, tag: str) -> Optional[Block]:
        if tag in ("latest", "pending", "safe", "finalized"):
            return self.blocks[-1]
        n = 0 if tag == "earliest" else int(tag, 16)
        for b in reversed(self.blocks):
            if b.number == n:
                return b
            if b.number < n:
                break
        # blocks skipped by hardhat_mine-style jumps: synthesise from the previous materialised block
        prev = [b for b in self.blocks if b.number < n]
        if prev and n <= self.blocks[-1].number:
            return Block(n, prev[-1].timestamp)
        return None
