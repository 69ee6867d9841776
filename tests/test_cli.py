"""Operator CLI (the reference's Hardhat tasks, contract/tasks/index.ts) against the devchain over HTTP."""
import json

import pytest

from arbius_b200 import cli
from arbius_b200.chain.devnode import DevChain
from arbius_b200.chain.engine_sim import UNIT
from arbius_b200.chain.eth import address_of
from arbius_b200.chain.rpc_server import serve_in_thread
from arbius_b200.protocol import template_bytes

DEPLOYER = address_of(cli.HARDHAT_KEY0)
KEY_B = "0x59c6995e998f97a5a0044966f0945389dc9e86dae88c7a8412f4603b6b78690d"


@pytest.fixture()
def net():
    chain = DevChain(deployer=DEPLOYER, funded={DEPLOYER: 40_000 * UNIT}, engine_balance=597000 * UNIT)
    srv, url = serve_in_thread(chain)
    yield chain, url
    srv.shutdown()
    srv.server_close()


def run(url, capsys, *argv, key=None, rc=0):
    args = ["--rpc-url", url] + (["--private-key", key] if key else []) + list(argv)
    assert cli.main(args) == rc
    return capsys.readouterr()


def test_task_list_matches_reference_names():
    ref = {"accounts", "decode-tx", "mine", "timetravel", "gen-wallet", "send-eth", "mining:balance", "mining:transfer",
           "mining:claimSolution", "mining:signalSupport", "model:register", "validator:stake", "validator:deploy-delegation",
           "admin:transferOwnership", "admin:setSolutionMineableStatus", "engine:isPaused", "engine:pause",
           "treasury:withdrawAccruedFees", "governance:delegate", "governance:propose", "governance:vote",
           "governance:cancel", "governance:queue", "governance:execute", "governance:proposal"}
    assert ref <= set(cli.TASKS)
    assert cli.parse_ether("1.000012") == 1000012 * 10 ** 12 and cli.format_ether(15 * UNIT // 10) == "1.5"
    assert cli.format_ether(3 * UNIT) == "3.0"


def test_mining_model_validator_and_admin_tasks(net, capsys, tmp_path):
    chain, url = net
    assert DEPLOYER[2:] in run(url, capsys, "accounts").out.lower()
    assert "balance: 40000.0" in run(url, capsys, "mining:balance").out
    tpl = tmp_path / "anythingv3.json"
    tpl.write_bytes(template_bytes("anythingv3"))
    out = run(url, capsys, "model:register", "--address", DEPLOYER, "--fee", "0", "--template", str(tpl)).out
    mid = out.split("Model registered: ")[1].split()[0]
    assert mid in chain.engine.models and '"mineable": false' in out
    run(url, capsys, "admin:setSolutionMineableStatus", "--model", mid, "--enabled", "true")
    assert chain.engine.models[mid].rate == UNIT
    run(url, capsys, "validator:stake")
    assert chain.engine.validator(DEPLOYER).staked == 24 * UNIT // 10 * 1200 // 1000
    run(url, capsys, "mining:signalSupport", "--model", mid, "--support", "true")
    run(url, capsys, "mining:transfer", "--address", address_of(KEY_B), "--amount", "2.5")
    assert chain.token.balance_of(address_of(KEY_B)) == 25 * UNIT // 10
    # non-owner cannot administer; error is reported, exit code 1
    err = run(url, capsys, "engine:pause", "--pause", "true", key=KEY_B, rc=1).err
    assert "not pauser" in err
    run(url, capsys, "engine:pause", "--pause", "true")
    assert "Engine is paused: True" in run(url, capsys, "engine:isPaused").out
    run(url, capsys, "engine:pause", "--pause", "false")
    # a task through the chain, decoded back by decode-tx; claimSolution after the delay
    from arbius_b200.chain import contract_abi as CA
    from arbius_b200.chain.jsonrpc import Contract, JsonRpcProvider
    eng = Contract(JsonRpcProvider(url), chain.engine.address, CA.ENGINE_FUNCTIONS, CA.ENGINE_EVENTS, cli.HARDHAT_KEY0)
    rc = eng.send("submitTask", 0, DEPLOYER, mid, 0, b'{"prompt":"cat","negative_prompt":""}')
    tid = eng.decode_logs(rc["logs"])[0].args["id"]
    dec = json.loads(run(url, capsys, "decode-tx", "--txid", rc["transactionHash"]).out)
    assert dec["input"]["prompt"] == "cat" and dec["model"] == mid
    cid = "0x1220" + "ab" * 32
    eng.send("signalCommitment", chain.engine.generate_commitment(DEPLOYER, tid, cid))
    eng.send("submitSolution", tid, cid)
    assert "not enough delay" in run(url, capsys, "mining:claimSolution", "--taskid", tid, rc=1).err
    run(url, capsys, "timetravel", "--seconds", "2100")
    run(url, capsys, "mine", "--blocks", "3")
    assert "Solution claimed" in run(url, capsys, "mining:claimSolution", "--taskid", tid).out
    run(url, capsys, "treasury:withdrawAccruedFees")
    run(url, capsys, "admin:transferOwnership", "--address", address_of(KEY_B))
    assert chain.engine.owner == address_of(KEY_B)
    out = run(url, capsys, "gen-wallet").out
    assert "privateKey 0x" in out
    assert "not deployed" in str(pytest.raises(SystemExit, cli.main, ["--rpc-url", url, "validator:deploy-delegation",
                                                                     "--modelIds", mid]).value)


def test_governance_tasks_and_deploy(net, capsys, tmp_path):
    chain, url = net
    cfg = tmp_path / "config.json"
    out = run(url, capsys, "deploy", "--out", str(cfg)).out
    conf = json.loads(cfg.read_text())
    mid = conf["models"]["kandinsky2"]["id"]
    assert conf["models"]["kandinsky2"]["params"]["cid"] == "0x1220511fdf0e88fa9adba98a7693cca89b5d9f3181da4815b0f0500aa4315b38d17d"
    assert chain.engine.models[mid].rate == UNIT and mid in out
    run(url, capsys, "admin:transferOwnership", "--address", chain.timelock.address)
    run(url, capsys, "governance:delegate", "--address", DEPLOYER)
    md = tmp_path / "proposal.md"
    md.write_text("# Lower the solution fee\n")
    actions = tmp_path / "actions.json"
    actions.write_text(json.dumps([{"target": chain.engine.address, "function": "setSolutionFeePercentage", "args": [5 * 10 ** 16]}]))
    out = run(url, capsys, "--config", str(cfg), "governance:propose", "--f", str(md), "--actions", str(actions)).out
    pid = out.split("Proposal submitted ")[1].split()[0]
    run(url, capsys, "mine", "--blocks", "6576")
    run(url, capsys, "governance:vote", "--proposal", pid, "--vote", "1")
    run(url, capsys, "mine", "--blocks", "6576")
    info = run(url, capsys, "--config", str(cfg), "governance:proposal", "--proposal", pid).out
    assert "state Succeeded" in info and "forVotes 40000.0" in info and "EngineV1.setSolutionFeePercentage(50000000000000000)" in info
    run(url, capsys, "governance:queue", "--proposal", pid)
    run(url, capsys, "governance:execute", "--proposal", pid)
    assert chain.engine.solution_fee_percentage == 5 * 10 ** 16
    assert "state Executed" in run(url, capsys, "governance:proposal", "--proposal", pid).out
    # a second proposal is cancelled by its proposer
    md.write_text("# Another one\n")
    out = run(url, capsys, "--config", str(cfg), "governance:propose", "--f", str(md), "--actions", str(actions)).out
    pid2 = out.split("Proposal submitted ")[1].split()[0]
    run(url, capsys, "governance:cancel", "--proposal", pid2)
    assert "state Canceled" in run(url, capsys, "governance:proposal", "--proposal", pid2).out
