"""Ethereum wire primitives, the devchain JSON-RPC node, the JSON-RPC chain client and governance.

The reference exercises its contracts on Hardhat's in-process EVM and its miner only against live chains
(SURVEY §4); here the same flows run over real HTTP JSON-RPC with signed raw transactions against DevChain.
"""
import json

import pytest

from arbius_b200.chain import contract_abi as CA
from arbius_b200.chain.client import ChainError
from arbius_b200.chain.devnode import DevChain
from arbius_b200.chain.engine_sim import UNIT
from arbius_b200.chain.eth import (Transaction, address_of, decode_raw_transaction, generate_wallet, recover, rlp_decode,
                                   rlp_encode, sign, sign_transaction, to_checksum_address)
from arbius_b200.chain.governance import (ACTIVE, EXECUTED, QUEUED, SUCCEEDED, GatewayMock, L1TokenSim, compute_pool_address,
                                          pool_key)
from arbius_b200.chain.jsonrpc import Contract, JsonRpcChainClient, JsonRpcProvider
from arbius_b200.chain.rpc_server import serve_in_thread
from arbius_b200.miner.db import MinerDB
from arbius_b200.miner.ml import NativeBackend
from arbius_b200.miner.node import MinerNode
from arbius_b200.protocol import MiningFilter, ModelEntry, abi, load_template, solution_cid, template_bytes
from arbius_b200.protocol.config import parse_mining_config
from arbius_b200.protocol.keccak import keccak256

KEY_DEPLOYER = "0xac0974bec39a17e36ba4a6b4d238ff944bacb478cbed5efcae784d7bf4f2ff80"     # Hardhat account #0
KEY_MINER = "0x59c6995e998f97a5a0044966f0945389dc9e86dae88c7a8412f4603b6b78690d"        # Hardhat account #1
KEY_USER = "0x5de4111afa1a4b94908f83103eb1f1706367c2e68ca870fc3fb9a804cdab365a"         # Hardhat account #2
DEPLOYER, MINER, USER = address_of(KEY_DEPLOYER), address_of(KEY_MINER), address_of(KEY_USER)
FIXED_PNG = b"\x89PNG\r\n\x1a\n" + b"fixed-output" * 10


def test_secp256k1_rlp_and_eip155_vectors():
    assert to_checksum_address(DEPLOYER) == "0xf39Fd6e51aad88F6F4ce6aB8827279cffFb92266"
    assert MINER == "0x70997970c51812dc3a010c7d01b50e0d17dc79c8"
    # EIP-155 worked example
    tx = Transaction(chain_id=1, nonce=9, to="0x" + "35" * 20, value=10 ** 18, gas=21000, gas_price=20 * 10 ** 9)
    assert tx.signing_hash().hex() == "daf5a779ae972f972197303d7b574746c7ef83eadac0f2791ad23db92e4c8e53"
    raw = sign_transaction(tx, "0x" + "46" * 32)
    assert raw.hex() == ("f86c098504a817c800825208943535353535353535353535353535353535353535880de0b6b3a76400008025a028ef61340bd939"
                         "bc2195fe537567866003e1a15d3c71ff63e1590620aa636276a067cbe9d8997f761aecb703304b3800ccf555c9f3dc64214b29"
                         "7fb1966a3b6d83")
    f = rlp_decode(raw)
    assert int.from_bytes(f[6], "big") == 37
    assert int.from_bytes(f[7], "big") == 18515461264373351373200002665853028612451056578545711640558177340181847433846
    assert int.from_bytes(f[8], "big") == 46948507304638947509940763649030358759909902576025900602547168820602576006531
    back = decode_raw_transaction(raw)
    assert back.sender == address_of("0x" + "46" * 32) and back.value == 10 ** 18 and back.nonce == 9
    # type-2 round trip and signature recovery
    tx2 = Transaction(chain_id=31337, nonce=3, to=MINER, data=b"\x01\x02", gas=10 ** 6, max_fee_per_gas=10 ** 9,
                      max_priority_fee_per_gas=2)
    d2 = decode_raw_transaction(sign_transaction(tx2, KEY_USER))
    assert d2.sender == USER and d2.is_1559 and d2.data == b"\x01\x02"
    h = keccak256(b"hello")
    rec, r, s = sign(KEY_MINER, h)
    assert "0x" + keccak256(recover(h, rec, r, s))[12:].hex() == MINER
    # RLP canonical cases
    assert rlp_encode(b"dog") == b"\x83dog" and rlp_encode([]) == b"\xc0" and rlp_encode(0) == b"\x80"
    assert rlp_encode(1024) == b"\x82\x04\x00" and rlp_encode([b"cat", b"dog"]) == b"\xc8\x83cat\x83dog"
    assert rlp_decode(rlp_encode([b"a" * 60, [b"", b"\x01"]])) == [b"a" * 60, [b"", b"\x01"]]
    k, a = generate_wallet(b"seed")
    assert address_of(k) == a


def test_abi_arrays_tuples_and_event_tables():
    types = ["uint256", "uint32[]", "bytes10", "bytes"]
    enc = abi.encode(types, [0x123, [0x456, 0x789], b"1234567890", b"Hello, world!"])
    assert enc.hex().endswith("48656c6c6f2c20776f726c642100000000000000000000000000000000000000") and len(enc) == 9 * 32
    assert abi.decode(types, enc) == [0x123, [0x456, 0x789], b"1234567890", b"Hello, world!"]
    nested = abi.encode(["uint256[][]", "string[]"], [[[1, 2], [3]], ["one", "two", "three"]])
    assert abi.decode(["uint256[][]", "string[]"], nested) == [[[1, 2], [3]], ["one", "two", "three"]]
    assert len(CA.ENGINE_FUNCTIONS) == 77 and len(CA.ENGINE_EVENTS) == 31
    # well-known selectors / topics
    assert CA.selector_of("transfer", CA.TOKEN_FUNCTIONS).hex() == "a9059cbb"
    assert CA.event_topic("Transfer", CA.TOKEN_EVENTS).hex() == "ddf252ad1be2c89b69c2b068fc378daa952ba7f163c4a11628f55a4df523b3ef"
    topics, data = CA.encode_log(CA.ENGINE_EVENTS, "TaskSubmitted", {"id": "0x" + "01" * 32, "model": "0x" + "02" * 32, "fee": 7,
                                                                    "sender": MINER})
    assert len(topics) == 4 and int(data, 16) == 7
    name, args = CA.decode_log(CA.ENGINE_EVENTS, topics, data)
    assert name == "TaskSubmitted" and args == {"id": "0x" + "01" * 32, "model": "0x" + "02" * 32, "fee": 7, "sender": MINER}


@pytest.fixture()
def node():
    # engine balance 597000 -> pseudo supply 3000, so validators must stake (as in the miner tests)
    chain = DevChain(deployer=DEPLOYER, funded={MINER: 50 * UNIT, USER: 50 * UNIT}, engine_balance=597000 * UNIT)
    srv, url = serve_in_thread(chain)
    yield chain, url
    srv.shutdown()
    srv.server_close()


def test_rpc_surface_and_revert_atomicity(node):
    chain, url = node
    p = JsonRpcProvider(url)
    assert p.chain_id() == 31337 and p.request("arbius_getAddresses")["engineAddress"] == chain.engine.address
    eng = Contract(p, chain.engine.address, CA.ENGINE_FUNCTIONS, CA.ENGINE_EVENTS, KEY_USER)
    tok = Contract(p, chain.token.address, CA.TOKEN_FUNCTIONS, CA.TOKEN_EVENTS, KEY_USER)
    assert tok.call("symbol") == "AIUS" and tok.call("balanceOf", USER) == 50 * UNIT
    assert eng.call("getPsuedoTotalSupply") == 3000 * UNIT and eng.call("getValidatorMinimum") == 24 * UNIT // 10
    assert eng.call("targetTs", 0) == 0 and eng.call("paused") is False
    rc = eng.send("registerModel", USER, 0, template_bytes("anythingv3"))
    mid = eng.decode_logs(rc["logs"])[0].args["id"]
    assert eng.call("models", mid)[1] == USER
    # hashModel over a tuple argument equals the registered id
    fee, addr, rate, cid = eng.call("models", mid)
    assert "0x" + eng.call("hashModel", (fee, addr, rate, cid), USER).hex() == mid
    # a reverted transaction leaves no trace: fee 1 AIUS without allowance
    block0, nlogs = p.block_number(), len(chain.logs)
    with pytest.raises(ChainError, match="insufficient allowance"):
        eng.send("submitTask", 0, USER, mid, UNIT, b'{"prompt":"x","negative_prompt":""}')
    assert p.block_number() == block0 + 1 and len(chain.logs) == nlogs and chain.engine.tasks == {}
    assert p.nonce(USER) == 2                                        # the nonce is still consumed
    with pytest.raises(ChainError, match="model does not exist"):
        eng.call("submitTask", 0, USER, "0x" + "00" * 32, 0, b"{}")
    # eth_call of a mutating function returns its value and rolls back
    tid = eng.call("submitTask", 0, USER, mid, 0, b'{"prompt":"x","negative_prompt":""}')
    assert len(tid) == 32 and chain.engine.tasks == {}
    # time travel + snapshots
    t0 = int(p.block("latest")["timestamp"], 16)
    snap = p.request("evm_snapshot")
    p.request("evm_increaseTime", [hex(5000)])
    p.request("hardhat_mine", [hex(10)])
    assert int(p.block("latest")["timestamp"], 16) >= t0 + 5000 and p.block_number() == block0 + 11
    assert p.request("evm_revert", [snap]) is True and p.block_number() == block0 + 1
    with pytest.raises(Exception, match="does not exist"):
        p.request("eth_noSuchMethod")


def test_miner_runs_against_jsonrpc_node(node):
    chain, url = node
    p = JsonRpcProvider(url)
    user_eng = Contract(p, chain.engine.address, CA.ENGINE_FUNCTIONS, CA.ENGINE_EVENTS, KEY_USER)
    user_tok = Contract(p, chain.token.address, CA.TOKEN_FUNCTIONS, CA.TOKEN_EVENTS, KEY_USER)
    rc = user_eng.send("registerModel", USER, 0, template_bytes("anythingv3"))
    mid = user_eng.decode_logs(rc["logs"])[0].args["id"]
    model = ModelEntry(id=mid, template=load_template("anythingv3"), filters=[MiningFilter()], name="anythingv3")
    cfg = parse_mining_config({"db_path": ":memory:", "stake_buffer_percent": 20, "stake_buffer_topup_percent": 1,
                               "ml": {"strategy": "native"}, "ipfs": {"strategy": "local"}, "automine": {"enabled": False}})
    client = JsonRpcChainClient(url, KEY_MINER, chain.engine.address, chain.token.address)
    assert client.address == MINER and client.token_balance() == 50 * UNIT
    calls = []
    ml = NativeBackend({mid: [lambda inp: (calls.append(inp), FIXED_PNG)[1]]})
    miner = MinerNode(cfg, client, ml, [model], db=MinerDB(":memory:"), n_gpus=1, sleep=lambda s: None)
    miner.bootup()
    miner.run_until_idle()
    assert client.validator_staked() == 3 * UNIT                     # min 2.4 with a 20 % buffer
    user_tok.send("approve", chain.engine.address, UNIT)
    rc = user_eng.send("submitTask", 0, USER, mid, UNIT, json.dumps({"prompt": "arbius test cat", "negative_prompt": ""},
                                                                  separators=(",", ":")).encode())
    tid = user_eng.decode_logs(rc["logs"])[0].args["id"]
    assert client.task_input_bytes(tid, rc["transactionHash"]).startswith(b'{"prompt"')
    miner.run_until_idle()
    sol = client.solutions(tid)
    assert sol.validator == MINER and sol.cid == solution_cid({"out-1.png": FIXED_PNG}) and len(calls) == 1
    names = [CA.decode_log(CA.ENGINE_EVENTS, lg["topics"], lg["data"])[0] for lg in chain.logs
             if lg["address"] == chain.engine.address]
    assert names.index("SignalCommitment") < names.index("SolutionSubmitted")
    before = client.token_balance()
    p.request("evm_increaseTime", [hex(2200)])
    p.request("evm_mine")
    miner.run_until_idle()
    assert client.solutions(tid).claimed and client.token_balance() == before + UNIT * 9 // 10
    assert miner.db.get_failed_jobs() == []


def test_governance_proposal_lifecycle_over_rpc(node):
    chain, url = node
    p = JsonRpcProvider(url)
    p.request("arbius_mint", [USER, hex(29_950 * UNIT)])                  # 30000 of ~627000 AIUS: above the 4 % quorum
    gov = Contract(p, chain.governor.address, CA.GOVERNOR_FUNCTIONS, CA.GOVERNOR_EVENTS, KEY_USER)
    tok = Contract(p, chain.token.address, CA.TOKEN_FUNCTIONS, CA.TOKEN_EVENTS, KEY_USER)
    eng_owner = Contract(p, chain.engine.address, CA.ENGINE_FUNCTIONS, CA.ENGINE_EVENTS, KEY_DEPLOYER)
    assert gov.call("votingDelay") == 6575 and gov.call("votingPeriod") == 6575 and gov.call("proposalThreshold") == UNIT
    rc = eng_owner.send("registerModel", USER, 0, template_bytes("kandinsky2"))
    mid = eng_owner.decode_logs(rc["logs"])[0].args["id"]
    eng_owner.send("transferOwnership", chain.timelock.address)            # governance owns the engine from here on
    with pytest.raises(ChainError, match="caller is not the owner"):
        eng_owner.send("setSolutionMineableRate", mid, UNIT)
    calldata = CA.encode_call(CA.ENGINE_FUNCTIONS, "setSolutionMineableRate", [mid, UNIT])
    desc = "# Make kandinsky2 mineable\nrate 1e18"
    with pytest.raises(ChainError, match="below proposal threshold"):
        gov.send("propose", [chain.engine.address], [0], [calldata], desc)
    tok.send("delegate", USER)
    assert tok.call("getVotes", USER) == 30_000 * UNIT
    rc = gov.send("propose", [chain.engine.address], [0], [calldata], desc)
    ev = gov.decode_logs(rc["logs"])[0]
    pid = ev.args["proposalId"]
    dh = keccak256(desc.encode())
    assert ev.name == "ProposalCreated" and pid == gov.call("hashProposal", [chain.engine.address], [0], [calldata], dh)
    assert gov.call("proposalsCreatedLength") == 1 and gov.call("proposalsCreated", 0) == pid
    assert gov.call("descriptionCids", pid) == bytes.fromhex(chain.engine.generate_ipfs_cid(desc.encode())[2:])
    assert gov.call("state", pid) == 0
    p.request("hardhat_mine", [hex(6576)])
    assert gov.call("state", pid) == ACTIVE
    gov.send("castVote", pid, 1)
    with pytest.raises(ChainError, match="vote already cast"):
        gov.send("castVote", pid, 1)
    assert gov.call("proposalVotes", pid) == [0, 30_000 * UNIT, 0]
    p.request("hardhat_mine", [hex(6576)])
    assert gov.call("state", pid) == SUCCEEDED                           # for-votes above the 4 % quorum
    gov.send("queue", [chain.engine.address], [0], [calldata], dh)
    assert gov.call("state", pid) == QUEUED
    gov.send("execute", [chain.engine.address], [0], [calldata], dh)
    assert gov.call("state", pid) == EXECUTED and eng_owner.call("models", mid)[2] == UNIT


def test_l1_token_gateway_handshake_and_pool_address():
    gw = GatewayMock()
    l1 = L1TokenSim(DEPLOYER, gw, gw, 1_000_000)
    assert l1.balance_of(DEPLOYER) == 1_000_000 * UNIT
    with pytest.raises(Exception, match="NOT_EXPECTED_CALL"):
        l1.is_arbitrum_enabled()
    with pytest.raises(Exception, match="not the owner"):
        l1.register_token_on_l2(USER, "0x" + "b0" * 20)
    l1.register_token_on_l2(DEPLOYER, "0x" + "b0" * 20)
    assert gw.l1_to_l2[l1.address] == "0x" + "b0" * 20 and gw.gateways[l1.address] is gw
    with pytest.raises(Exception, match="NOT_EXPECTED_CALL"):
        l1.is_arbitrum_enabled()
    # Uniswap v3 USDC/WETH 0.05 % pool on mainnet
    key = pool_key("0xC02aaA39b223FE8D0A0e5C4F27eAD9083C756Cc2", "0xA0b86991c6218b36c1d19D4a2e9Eb0cE3606eB48", 500)
    assert key[0] == "0xa0b86991c6218b36c1d19d4a2e9eb0ce3606eb48"
    assert compute_pool_address("0x1F98431c8aD98523631AE4a59f267346ea31F984", key) == "0x88e6a0c2ddd26feeb64f039a2c41296fcb3f5640"
