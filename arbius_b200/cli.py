"""Operator CLI: the reference's Hardhat tasks as ``python -m arbius_b200.cli <task> [--param value]``.

Same task names and parameters as contract/tasks/index.ts:12-465 (``accounts``, ``decode-tx``, ``mine``, ``timetravel``,
``gen-wallet``, ``send-eth``, ``mining:*``, ``model:register``, ``validator:stake``, ``validator:deploy-delegation``,
``admin:*``, ``engine:*``, ``treasury:withdrawAccruedFees``, ``governance:*``) plus ``node`` (start the devchain, the
``npx hardhat node`` stand-in) and ``deploy`` (what scripts/003-deploy-core-basic.ts leaves behind: Kandinsky2 registered
as a free mineable model, addresses written to a config file).

Network selection follows contract/utils/config.ts:34-80: ``--network nova`` reads ``NOVA_PROVIDER_URL`` /
``NOVA_PRIVATE_KEY`` from the environment; ``--rpc-url`` / ``--private-key`` override; the default network is
``localhost`` (http://127.0.0.1:8545, Hardhat account #0). Contract addresses come from ``--config`` (a
scripts/config.json-shaped file), else from the devchain's ``arbius_getAddresses``, else from the Nova deployment.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from typing import Any, Callable, Dict, List, Optional

from .chain import contract_abi as CA
from .chain.client import ChainError
from .chain.eth import Transaction, address_of, generate_wallet, sign_transaction, to_checksum_address
from .chain.jsonrpc import Contract, JsonRpcError, JsonRpcProvider
from .protocol import abi
from .protocol.keccak import keccak256

NETWORKS = ("localhost", "mainnet", "goerli", "sepolia", "arbgoerli", "arbsepolia", "nova")
HARDHAT_KEY0 = "0xac0974bec39a17e36ba4a6b4d238ff944bacb478cbed5efcae784d7bf4f2ff80"
UNIT = 10 ** 18


def get_env(key: str, default: Any = None) -> Any:
    """reference: contract/utils/config.ts:34-50 — required unless a default is given, 'true'/'false' coerced"""
    v = os.environ.get(key)
    if not v:
        if default is None:
            raise KeyError(f"Required env var {key} not set")
        return default
    if v.lower() in ("true", "false"):
        return v.lower() == "true"
    return v


def parse_ether(s: str) -> int:
    whole, _, frac = str(s).partition(".")
    return int(whole or "0") * UNIT + int((frac + "0" * 18)[:18] or "0")


def format_ether(v: int) -> str:
    s = f"{v // UNIT}.{v % UNIT:018d}".rstrip("0")
    return s + "0" if s.endswith(".") else s


class Env:
    """provider + signer + attached contracts for one CLI invocation (the ``hre`` of a Hardhat task)"""

    def __init__(self, a: argparse.Namespace):
        net = a.network
        if a.rpc_url:
            url = a.rpc_url
        elif net == "localhost":
            url = "http://127.0.0.1:8545"
        else:
            url = get_env(f"{net.upper()}_PROVIDER_URL")
        key = a.private_key or (HARDHAT_KEY0 if net == "localhost" else get_env(f"{net.upper()}_PRIVATE_KEY"))
        self.key = key if key.startswith("0x") else "0x" + key
        self.provider = JsonRpcProvider(url)
        self.address = address_of(self.key)
        cfg: Dict[str, Any] = {}
        if a.config and os.path.exists(a.config):
            with open(a.config) as f:
                cfg = json.load(f)
        else:
            try:
                cfg = self.provider.request("arbius_getAddresses")
            except (JsonRpcError, OSError):
                from .protocol.config import STATIC_CHAIN_CONFIG
                cfg = dict(STATIC_CHAIN_CONFIG)
        self.config = cfg
        self.out: List[str] = []

    def _contract(self, key: str, fns, evs) -> Contract:
        addr = self.config.get(key)
        if not addr:
            raise SystemExit(f"error: {key} is not configured for this network")
        return Contract(self.provider, addr, fns, evs, self.key)

    @property
    def engine(self) -> Contract: return self._contract("engineAddress", CA.ENGINE_FUNCTIONS, CA.ENGINE_EVENTS)
    @property
    def token(self) -> Contract: return self._contract("baseTokenAddress", CA.TOKEN_FUNCTIONS, CA.TOKEN_EVENTS)
    @property
    def governor(self) -> Contract: return self._contract("governorAddress", CA.GOVERNOR_FUNCTIONS, CA.GOVERNOR_EVENTS)

    def log(self, *parts) -> None:
        line = " ".join(str(p) for p in parts)
        self.out.append(line)
        print(line)


TASKS: Dict[str, dict] = {}


def task(name: str, desc: str, params: Optional[List[tuple]] = None):
    def deco(fn: Callable[[Env, argparse.Namespace], None]):
        TASKS[name] = {"fn": fn, "desc": desc, "params": params or []}
        return fn
    return deco


def _bool(s: str) -> bool:
    return str(s).lower() in ("1", "true", "yes", "y")


def _hex32(b) -> str:
    return "0x" + bytes(b).hex() if isinstance(b, (bytes, bytearray)) else b


# ---------------------------------------------------------------------------------------------- basics
@task("accounts", "Prints the list of accounts")
def _accounts(env, a):
    env.log(to_checksum_address(env.address))


@task("decode-tx", "Extract input from a submitTask transaction", [("txid", "transaction hash")])
def _decode_tx(env, a):
    tx = env.provider.request("eth_getTransactionByHash", [a.txid])
    if tx is None:
        raise SystemExit("error: transaction not found")
    sig = CA.signature("submitTask", CA.ENGINE_FUNCTIONS["submitTask"][0])
    version, owner, model, fee, inp = abi.decode_call(sig, bytes.fromhex(tx["input"][2:]))
    env.log(json.dumps({"version": version, "owner": owner, "model": _hex32(model), "fee": str(fee),
                        "input": json.loads(inp.decode())}, indent=2))


@task("mine", "mine blocks locally", [("blocks", "how many")])
def _mine(env, a):
    env.provider.request("hardhat_mine", [hex(int(a.blocks))])
    env.log(f"mined {int(a.blocks)} blocks, now at {env.provider.block_number()}")


@task("timetravel", "go into the future", [("seconds", "how many")])
def _timetravel(env, a):
    env.provider.request("evm_increaseTime", [hex(int(a.seconds))])
    env.provider.request("evm_mine", [])
    env.log(f"time is now {int(env.provider.block('latest')['timestamp'], 16)}")


@task("gen-wallet", "Creates a new wallet")
def _gen_wallet(env, a):
    key, addr = generate_wallet()
    env.log("address", to_checksum_address(addr))
    env.log("privateKey", key)


@task("send-eth", "Send ether", [("address", "Receiver address"), ("amount", "Eth")])
def _send_eth(env, a):
    p = env.provider
    tx = Transaction(chain_id=p.chain_id(), nonce=p.nonce(env.address), to=a.address.lower(), value=parse_ether(a.amount),
                     gas=21000, gas_price=p.gas_price())
    h = p.request("eth_sendRawTransaction", ["0x" + sign_transaction(tx, env.key).hex()])
    env.log(f"sent {a.amount} eth to {a.address}: {h}")


# ---------------------------------------------------------------------------------------------- mining
@task("mining:balance", "Get miners token balance")
def _balance(env, a):
    env.log(f"balance: {format_ether(env.token.call('balanceOf', env.address))}")


@task("mining:transfer", "Transfer mining rewards to new address", [("address", "Receiver address"), ("amount", "In token value")])
def _transfer(env, a):
    env.token.send("transfer", a.address, parse_ether(a.amount), gas=100_000)
    env.log(f"Transferred {a.amount} to {a.address}")


@task("mining:claimSolution", "Claim past task", [("taskid", "Task id")])
def _claim(env, a):
    rc = env.engine.send("claimSolution", a.taskid, gas=300_000)
    env.log(f"Solution claimed in {rc['transactionHash']}")


@task("mining:signalSupport", "Signal mining support for model", [("model", "Model id"), ("support", "Whether supported or not")])
def _signal_support(env, a):
    env.engine.send("signalSupport", a.model, _bool(a.support))
    env.log(f"Signalled support={_bool(a.support)} for {a.model}")


@task("model:register", "Register new model", [("address", "Address for fees for model"), ("fee", "Flat fee to set"),
                                               ("template", "Model schema")])
def _model_register(env, a):
    with open(a.template, "rb") as f:
        tdata = f.read()
    if len(tdata) > 262144:
        raise SystemExit("error: template file bigger than 262144 bytes")
    rc = env.engine.send("registerModel", a.address, int(a.fee), tdata, gas=3_000_000)
    mid = env.engine.decode_logs(rc["logs"])[0].args["id"]
    cid = env.engine.call("models", mid)[3]
    env.log(f"Model registered: {mid}")
    env.log()
    env.log('Add the following to the "models" section of your config:')
    env.log()
    env.log(json.dumps({"<your-model-name>": {"id": mid, "mineable": False, "contracts": {},
                                              "params": {"addr": a.address, "fee": str(a.fee), "cid": _hex32(cid)}}}, indent=2))


@task("validator:stake", "Become a validator")
def _validator_stake(env, a):
    minimum = env.engine.call("getValidatorMinimum")
    amount = minimum * 1200 // 1000
    if env.token.call("allowance", env.address, env.engine.address) < amount:
        env.token.send("approve", env.engine.address, (1 << 256) - 1, gas=100_000)
    env.engine.send("validatorDeposit", env.address, amount, gas=500_000)
    env.log(f"{to_checksum_address(env.address)} is now a validator")


@task("validator:deploy-delegation", "Create delegated validator", [("modelIds", "model ids (comma separated)")])
def _deploy_delegation(env, a):
    # the reference task calls a DelegatedValidatorDeployerV1 contract that is not part of its tree
    raise SystemExit("error: DelegatedValidatorDeployerV1 is not deployed on this network (the reference tree has no such contract)")


# ---------------------------------------------------------------------------------------------- admin
@task("admin:transferOwnership", "Transfer admin ownership of Engine", [("address", "To who?")])
def _admin_owner(env, a):
    env.engine.send("transferOwnership", a.address)
    env.log(f"Transferred ownership to {a.address}")


@task("admin:setSolutionMineableStatus", "Allow a model to receive rewards for tasks completed",
      [("model", "Model id"), ("enabled", "Enable/Disable")])
def _admin_mineable(env, a):
    # the deployed contract names this setSolutionMineableRate (EngineV1.sol:294); enabled = rate 1e18
    env.engine.send("setSolutionMineableRate", a.model, UNIT if _bool(a.enabled) else 0)
    env.log(f"Model {a.model} mineable: {_bool(a.enabled)}")


@task("engine:isPaused", "Check if engine is paused")
def _is_paused(env, a):
    env.log(f"Engine is paused: {env.engine.call('paused')}")


@task("engine:pause", "Pause engine", [("pause", "Pause/Unpause")])
def _pause(env, a):
    env.engine.send("setPaused", _bool(a.pause))
    env.log(f"Engine is now {'paused' if env.engine.call('paused') else 'unpaused'}")


@task("treasury:withdrawAccruedFees", "Withdraw fees to treasury")
def _withdraw_fees(env, a):
    env.engine.send("withdrawAccruedFees")
    env.log("Fees withdrawn")


# ---------------------------------------------------------------------------------------------- governance
@task("governance:delegate", "Delegate your tokens", [("address", "To who? (can be yourself)")])
def _gov_delegate(env, a):
    env.token.send("delegate", a.address, gas=200_000)
    env.log(f"Delegated to {a.address}")


@task("governance:propose", "Propose a vote", [("f", "Filename of proposal markdown file"),
                                               ("actions", "optional JSON file: [{target, value, function, args}]", None)])
def _gov_propose(env, a):
    with open(a.f, encoding="utf8") as f:
        description = f.read()
    targets, values, calldatas = [], [], []
    if a.actions:
        tables = {env.config.get("engineAddress", "").lower(): CA.ENGINE_FUNCTIONS,
                  env.config.get("baseTokenAddress", "").lower(): CA.TOKEN_FUNCTIONS}
        with open(a.actions) as f:
            for act in json.load(f):
                table = tables[act["target"].lower()]
                targets.append(act["target"].lower())
                values.append(int(act.get("value", 0)))
                calldatas.append(CA.encode_call(table, act["function"], act["args"]))
    else:
        # the reference task hard-codes one example action: transfer 1.000012 AIUS from the timelock
        targets, values = [env.token.address], [0]
        calldatas = [CA.encode_call(CA.TOKEN_FUNCTIONS, "transfer",
                                    ["0x1a320e53a25f518b893f286f3600cc204c181a8e", parse_ether("1.000012")])]
    env.governor.send("propose", targets, values, calldatas, description, gas=3_000_000)
    pid = env.governor.call("hashProposal", targets, values, calldatas, keccak256(description.encode()))
    env.log("Proposal submitted", hex(pid))


def _actions(env, pid: int):
    targets, values, _sigs, calldatas = env.governor.call("getActions", pid)
    return targets, values, calldatas, env.governor.call("descriptionHashes", pid)


@task("governance:vote", "Vote on a proposal", [("proposal", "Proposal id"), ("vote", "0=against, 1=for, 2=abstain")])
def _gov_vote(env, a):
    env.governor.send("castVote", int(a.proposal, 0), int(a.vote))
    env.log("Vote cast")


@task("governance:cancel", "Cancel a proposal", [("proposal", "Proposal id")])
def _gov_cancel(env, a):
    env.governor.send("cancel", *_actions(env, int(a.proposal, 0)))
    env.log("Proposal cancelled")


@task("governance:queue", "Queue a proposal", [("proposal", "Proposal id")])
def _gov_queue(env, a):
    env.governor.send("queue", *_actions(env, int(a.proposal, 0)))
    env.log("Proposal queued")


@task("governance:execute", "Execute a proposal", [("proposal", "Proposal id")])
def _gov_execute(env, a):
    env.governor.send("execute", *_actions(env, int(a.proposal, 0)), gas=5_000_000)
    env.log("Proposal executed")


@task("governance:proposal", "Look up proposal info", [("proposal", "Proposal id")])
def _gov_proposal(env, a):
    pid = int(a.proposal, 0)
    (_id, proposer, eta, start, end, for_, against, abstain, canceled, executed) = env.governor.call("proposals", pid)
    from .chain.governance import STATE_NAMES
    for k, v in (("id", hex(pid)), ("proposer", proposer), ("eta", eta), ("startBlock", start), ("endBlock", end),
                 ("forVotes", format_ether(for_)), ("againstVotes", format_ether(against)),
                 ("abstainVotes", format_ether(abstain)), ("canceled", canceled), ("executed", executed),
                 ("state", STATE_NAMES[env.governor.call("state", pid)])):
        env.log(k, v)
    targets, values, calldatas, _dh = _actions(env, pid)
    tables = {env.config.get("engineAddress", "").lower(): ("EngineV1", CA.ENGINE_FUNCTIONS),
              env.config.get("baseTokenAddress", "").lower(): ("BaseTokenV1", CA.TOKEN_FUNCTIONS)}
    for i, (t, v, c) in enumerate(zip(targets, values, calldatas)):
        env.log()
        env.log(f"target {i}", t, "value", format_ether(v))
        ent = tables.get(t.lower())
        name = CA.selector_table(ent[1]).get(bytes(c[:4])) if ent else None
        if name:
            args = abi.decode(ent[1][name][0], bytes(c[4:]))
            env.log(f"{ent[0]}.{name}({', '.join(_hex32(x) if isinstance(x, (bytes, bytearray)) else str(x) for x in args)})")
        else:
            env.log("calldata", "0x" + bytes(c).hex())


# ---------------------------------------------------------------------------------------------- devchain
@task("deploy", "Register Kandinsky2 as a free mineable model and write the address/model config",
      [("out", "config file to write", "arbius_config.json")])
def _deploy(env, a):
    from .protocol import template_bytes
    models = {}
    for name in ("kandinsky2",):
        rc = env.engine.send("registerModel", "0x" + "00" * 19 + "01", 0, template_bytes(name), gas=3_000_000)
        mid = env.engine.decode_logs(rc["logs"])[0].args["id"]
        env.engine.send("setSolutionMineableRate", mid, UNIT)
        fee, addr, rate, cid = env.engine.call("models", mid)
        models[name] = {"id": mid, "mineable": True, "contracts": {},
                        "params": {"addr": addr, "fee": str(fee), "rate": str(rate), "cid": _hex32(cid)}}
        env.log(f"{name} registered: {mid}")
    cfg = dict(env.config)
    cfg["models"] = models
    with open(a.out, "w") as f:
        json.dump(cfg, f, indent=2)
    env.log(f"wrote {a.out}")


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="arbius_b200.cli", description=__doc__.split("\n\n")[0])
    ap.add_argument("--network", default="localhost", choices=NETWORKS)
    ap.add_argument("--rpc-url")
    ap.add_argument("--private-key")
    ap.add_argument("--config", help="scripts/config.json-shaped address file")
    sub = ap.add_subparsers(dest="task", required=True)
    for name, t in TASKS.items():
        sp = sub.add_parser(name, help=t["desc"])
        for prm in t["params"]:
            if len(prm) == 3:
                sp.add_argument(f"--{prm[0]}", help=prm[1], default=prm[2])
            else:
                sp.add_argument(f"--{prm[0]}", help=prm[1], required=True)
    node = sub.add_parser("node", help="start the devchain JSON-RPC node (npx hardhat node stand-in)")
    node.add_argument("rest", nargs=argparse.REMAINDER)
    gold = sub.add_parser("golden", help="boot self-test goldens: check them, or record them with --write "
                                          "(miner/src/index.ts:983-1001 keeps one hard-coded CID; here one per "
                                          "template, library build and GPU model)")
    gold.add_argument("mining_config", help="MiningConfig.json of the miner")
    gold.add_argument("--write", action="store_true", help="record the CIDs this box produces as the goldens")
    return ap


def golden_main(a) -> int:
    """Builds the miner exactly as ``python -m arbius_b200.miner`` would (same predictors, same IPFS strategy), runs
    the fixed probe on every enabled model and compares with / writes the golden file."""
    from .miner.node import FatalMinerError
    from .miner.start import start
    node = start(a.mining_config, run=False)
    try:
        got = node.boot_self_test(write=a.write)
    except FatalMinerError:
        print("golden mismatch: this box does not reproduce the recorded CIDs", file=sys.stderr)
        return 1
    for k, v in sorted(got.items()):
        print(("wrote   " if a.write else "matches ") + f"{k}  {v}")
    if not got:
        print("no goldens recorded for this (template, library, GPU); run with --write on a verified box")
    return 0


def main(argv: Optional[List[str]] = None) -> int:
    a = build_parser().parse_args(argv)
    if a.task == "node":
        from .chain.rpc_server import main as node_main
        return node_main(a.rest)
    if a.task == "golden":
        return golden_main(a)
    try:
        env = Env(a)
        TASKS[a.task]["fn"](env, a)
    except ChainError as e:
        print(f"error: {e.reason}", file=sys.stderr)
        return 1
    except KeyError as e:
        print(f"error: {e.args[0]}", file=sys.stderr)
        return 1
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
