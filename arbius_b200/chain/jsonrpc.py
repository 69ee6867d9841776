"""ChainClient over Ethereum JSON-RPC with locally signed transactions.

Plays the role of ethers' JsonRpcProvider + Wallet + Contract objects in the reference miner
(reference: miner/src/blockchain.ts:22-75): views are ``eth_call``, transactions are signed here (legacy EIP-155,
explicit gas limits as in miner/src/index.ts:483, 621, 636, 737) and sent with ``eth_sendRawTransaction``, events are
pulled with ``eth_getLogs`` over the block range since the last poll and decoded against the contract event tables.
"""
from __future__ import annotations

import itertools
import json
import threading
import time
import urllib.request
from typing import Any, Callable, Dict, List, Optional, Sequence

from ..protocol import abi
from . import contract_abi as CA
from .client import (GAS_LIMITS, ChainClient, ChainError, ContestationView, Receipt, SolutionView, TaskView)
from .engine_sim import Event
from .eth import Transaction, address_of, sign_transaction


class JsonRpcError(Exception):
    def __init__(self, code: int, message: str, data: Any = None):
        super().__init__(f"{message} (code {code})")
        self.code, self.message, self.data = code, message, data


def decode_revert(data: Optional[str]) -> Optional[str]:
    """Error(string) payload -> reason"""
    if not data or not data.startswith("0x08c379a0"):
        return None
    try:
        return abi.decode(["string"], bytes.fromhex(data[10:]))[0]
    except (ValueError, UnicodeDecodeError):
        return None


class JsonRpcProvider:
    def __init__(self, url: str, timeout: float = 30.0):
        self.url, self.timeout = url, timeout
        self._ids = itertools.count(1)

    def request(self, method: str, params: Sequence[Any] = ()) -> Any:
        body = json.dumps({"jsonrpc": "2.0", "id": next(self._ids), "method": method, "params": list(params)}).encode()
        req = urllib.request.Request(self.url, data=body, headers={"Content-Type": "application/json"})
        with urllib.request.urlopen(req, timeout=self.timeout) as resp:
            out = json.loads(resp.read())
        if "error" in out and out["error"]:
            e = out["error"]
            raise JsonRpcError(e.get("code", -32000), e.get("message", "rpc error"), e.get("data"))
        return out.get("result")

    # typed helpers
    def chain_id(self) -> int: return int(self.request("eth_chainId"), 16)
    def block_number(self) -> int: return int(self.request("eth_blockNumber"), 16)
    def gas_price(self) -> int: return int(self.request("eth_gasPrice"), 16)
    def balance(self, addr: str) -> int: return int(self.request("eth_getBalance", [addr, "latest"]), 16)
    def nonce(self, addr: str) -> int: return int(self.request("eth_getTransactionCount", [addr, "pending"]), 16)

    def block(self, tag="latest") -> dict:
        return self.request("eth_getBlockByNumber", [tag if isinstance(tag, str) else hex(tag), False])

    def call(self, to: str, data: bytes, sender: Optional[str] = None) -> bytes:
        tx = {"to": to, "data": "0x" + data.hex()}
        if sender:
            tx["from"] = sender
        res = self.request("eth_call", [tx, "latest"])
        return bytes.fromhex(res[2:])


# One lock and one local nonce counter per (provider, sender): the miner sends from several threads (one solve per GPU,
# plus claims / stake / automine on the main thread) through several Contract objects that share a key. Reading
# eth_getTransactionCount and sending must be atomic per sender, or two threads sign the same nonce and one
# transaction — a finished solve's commitment, say — is rejected ("nonce too low" / "replacement underpriced").
_sender_locks: Dict[tuple, threading.Lock] = {}
_sender_nonce: Dict[tuple, int] = {}
_sender_guard = threading.Lock()


def _sender_lock(key: tuple) -> threading.Lock:
    with _sender_guard:
        lk = _sender_locks.get(key)
        if lk is None:
            lk = _sender_locks[key] = threading.Lock()
        return lk


class Contract:
    """address + function/event tables bound to a provider (and optionally a signing key)"""

    def __init__(self, provider: JsonRpcProvider, address: str, functions: dict, events: dict,
                 private_key: Optional[str] = None, chain_id: Optional[int] = None):
        self.p, self.address, self.functions, self.events = provider, address.lower(), functions, events
        self.key = private_key
        self.sender = address_of(private_key) if private_key else None
        self._chain_id = chain_id

    def call(self, name: str, *args) -> Any:
        try:
            raw = self.p.call(self.address, CA.encode_call(self.functions, name, args), self.sender)
        except JsonRpcError as e:
            raise ChainError(decode_revert(e.data) or e.message) from e
        out = CA.decode_result(self.functions, name, raw)
        return out[0] if len(out) == 1 else out

    def send(self, name: str, *args, gas: int = 1_000_000, wait: bool = True, timeout: float = 120.0) -> dict:
        if self.key is None:
            raise ChainError("no private key configured for sending transactions")
        if self._chain_id is None:
            self._chain_id = self.p.chain_id()
        data = CA.encode_call(self.functions, name, args)
        skey = (id(self.p), self.sender)
        with _sender_lock(skey):
            for attempt in range(3):
                nonce = max(self.p.nonce(self.sender), _sender_nonce.get(skey, 0))
                tx = Transaction(chain_id=self._chain_id, nonce=nonce, to=self.address, data=data, gas=gas,
                                 gas_price=self.p.gas_price())
                try:
                    txhash = self.p.request("eth_sendRawTransaction", ["0x" + sign_transaction(tx, self.key).hex()])
                    _sender_nonce[skey] = nonce + 1
                    break
                except JsonRpcError as e:
                    msg = (decode_revert(e.data) or e.message or "")
                    if attempt < 2 and ("nonce" in msg.lower() or "underpriced" in msg.lower()):
                        _sender_nonce.pop(skey, None)            # another process used this key: resynchronise
                        continue
                    raise ChainError(msg) from e
        if not wait:
            return {"transactionHash": txhash}
        deadline = time.time() + timeout
        while True:
            rc = self.p.request("eth_getTransactionReceipt", [txhash])
            if rc is not None:
                break
            if time.time() > deadline:
                raise ChainError(f"timeout waiting for {txhash}")
            time.sleep(0.25)
        if int(rc.get("status", "0x1"), 16) == 0:
            reason = rc.get("revertReason")
            if reason is None:                                   # real nodes: replay as a call to learn the reason
                try:
                    self.p.call(self.address, data, self.sender)
                except JsonRpcError as e:
                    reason = decode_revert(e.data) or e.message
            raise ChainError(reason or "transaction reverted")
        return rc

    def decode_logs(self, logs: List[dict]) -> List[Event]:
        out = []
        for lg in logs:
            if lg["address"].lower() != self.address:
                continue
            dec = CA.decode_log(self.events, lg["topics"], lg["data"])
            if dec is not None:
                out.append(Event(dec[0], dec[1], int(lg["blockNumber"], 16), lg["transactionHash"]))
        return out


class JsonRpcChainClient(ChainClient):
    def __init__(self, rpc_url: str, private_key: str, engine_address: str, token_address: str,
                 poll_from_block: Optional[int] = None):
        self.provider = JsonRpcProvider(rpc_url)
        cid = self.provider.chain_id()
        self.engine = Contract(self.provider, engine_address, CA.ENGINE_FUNCTIONS, CA.ENGINE_EVENTS, private_key, cid)
        self.token = Contract(self.provider, token_address, CA.TOKEN_FUNCTIONS, CA.TOKEN_EVENTS, private_key, cid)
        self.address = self.engine.sender
        self.solver_address = self.engine.address
        self._handlers: List[Callable[[Event], None]] = []
        self._next_block = self.provider.block_number() + 1 if poll_from_block is None else poll_from_block

    # views
    def version(self): return self.engine.call("version")
    def now(self): return int(self.provider.block("latest")["timestamp"], 16)
    def block_number(self): return self.provider.block_number()
    def eth_balance(self): return self.provider.balance(self.address)

    @staticmethod
    def _h(b) -> str:
        return "0x" + bytes(b).hex() if isinstance(b, (bytes, bytearray)) else b

    def tasks(self, taskid):
        m, fee, owner, bt, ver, cid = self.engine.call("tasks", taskid)
        return TaskView(self._h(m), fee, owner, bt, ver, self._h(cid))

    def solutions(self, taskid):
        v, bt, claimed, cid = self.engine.call("solutions", taskid)
        return SolutionView(v, bt, claimed, self._h(cid))

    def contestations(self, taskid):
        v, bt, idx, slash = self.engine.call("contestations", taskid)
        return ContestationView(v, bt, idx, slash)

    def contestation_voted(self, taskid, addr): return self.engine.call("contestationVoted", taskid, addr)
    def validator_staked(self, addr=None): return self.engine.call("validators", addr or self.address)[0]
    def get_validator_minimum(self): return self.engine.call("getValidatorMinimum")
    def token_balance(self, addr=None): return self.token.call("balanceOf", addr or self.address)
    def token_allowance(self, owner, spender): return self.token.call("allowance", owner, spender)

    def task_input_bytes(self, taskid, txid):
        """reference: miner/src/index.ts:151-159 — decode submitTask calldata of the emitting transaction"""
        tx = self.provider.request("eth_getTransactionByHash", [txid])
        if tx is None:
            raise ChainError(f"transaction {txid} not found")
        data = bytes.fromhex(tx["input"][2:])
        sig = CA.signature("submitTask", CA.ENGINE_FUNCTIONS["submitTask"][0])
        return abi.decode_call(sig, data)[4]

    # transactions
    def _receipt(self, rc: dict) -> Receipt:
        return Receipt(rc["transactionHash"], int(rc["blockNumber"], 16))

    def token_approve(self, spender, amount): return self._receipt(self.token.send("approve", spender, amount, gas=100_000))

    def validator_deposit(self, validator, amount):
        return self._receipt(self.engine.send("validatorDeposit", validator, amount, gas=500_000))

    def submit_task(self, version, owner, model, fee, input_):
        rc = self.engine.send("submitTask", version, owner, model, fee, input_, gas=GAS_LIMITS["submitTask"])
        for ev in self.engine.decode_logs(rc.get("logs", [])):
            if ev.name == "TaskSubmitted":
                return ev.args["id"]
        raise ChainError("submitTask mined without a TaskSubmitted event")

    def signal_commitment(self, commitment):
        return self._receipt(self.engine.send("signalCommitment", commitment, gas=GAS_LIMITS["signalCommitment"]))

    def submit_solution(self, taskid, cid):
        return self._receipt(self.engine.send("submitSolution", taskid, cid, gas=GAS_LIMITS["submitSolution"]))

    def claim_solution(self, taskid):
        return self._receipt(self.engine.send("claimSolution", taskid, gas=GAS_LIMITS["claimSolution"]))

    def submit_contestation(self, taskid): return self._receipt(self.engine.send("submitContestation", taskid))
    def vote_on_contestation(self, taskid, yea): return self._receipt(self.engine.send("voteOnContestation", taskid, yea))

    def contestation_vote_finish(self, taskid, amnt):
        return self._receipt(self.engine.send("contestationVoteFinish", taskid, amnt, gas=3_000_000))

    # events
    def subscribe(self, handler): self._handlers.append(handler)

    def poll_events(self) -> int:
        head = self.provider.block_number()
        if head < self._next_block:
            return 0
        logs = self.provider.request("eth_getLogs", [{"address": self.engine.address, "fromBlock": hex(self._next_block),
                                                      "toBlock": hex(head)}])
        self._next_block = head + 1
        events = self.engine.decode_logs(logs)
        for ev in events:
            for h in self._handlers:
                h(ev)
        return len(events)
