"""Round-2 control-plane fixes: boot self-test with per-environment goldens (reference: miner/src/index.ts:983-1001),
JSON.parse-strict task inputs, automatic replay of failed solves, per-sender nonce serialisation, thread-local
canonical-batch multiplier, worker exit on device faults."""
import json
import sys
import threading
import time

import pytest

from arbius_b200.chain.client import SimChainClient
from arbius_b200.chain.engine_sim import UNIT, deploy_local
from arbius_b200.miner.db import MinerDB
from arbius_b200.miner.golden import GoldenStore, environment_key, selftest_input
from arbius_b200.miner.ml import NativeBackend
from arbius_b200.miner.node import FatalMinerError, MinerNode
from arbius_b200.protocol import MiningFilter, ModelEntry, load_template, solution_cid, template_bytes
from arbius_b200.protocol.config import parse_mining_config

MINER, USER = "0x" + "a1" * 20, "0x" + "11" * 20
PNG = b"\x89PNG\r\n\x1a\n" + b"golden" * 20


def _node(tmp_path, predict, goldens=None, **kw):
    eng = deploy_local(engine_balance=599990 * UNIT, funded={MINER: 50 * UNIT, USER: 50 * UNIT})
    mid = eng.register_model(USER, USER, 0, template_bytes("kandinsky2"))
    model = ModelEntry(id=mid, template=load_template("kandinsky2"), filters=[MiningFilter()], name="kandinsky2")
    cfg = parse_mining_config({"db_path": ":memory:", "ml": {"strategy": "native"}, "ipfs": {"strategy": "local"},
                               "automine": {"enabled": False}})
    node = MinerNode(cfg, SimChainClient(eng, MINER), NativeBackend({mid: [predict]}), [model], db=MinerDB(":memory:"),
                     sleep=lambda s: None, goldens=goldens, **kw)
    return eng, node, mid


def test_selftest_probe_is_the_reference_prompt_and_seed_plus_template_defaults():
    inp = selftest_input(load_template("kandinsky2"))
    assert inp["prompt"] == "arbius test cat" and inp["seed"] == 1337 and inp["width"] == 768 and inp["height"] == 768
    rvm = selftest_input(load_template("robust_video_matting"))
    assert rvm["input_video"] and rvm["seed"] == 1337


def test_boot_records_then_checks_goldens_and_a_flipped_golden_aborts_the_boot(tmp_path):
    seen = []
    store = GoldenStore(tmp_path / "goldens.json")
    _, node, _ = _node(tmp_path, lambda inp: (seen.append(inp), PNG)[1], goldens=store)
    node.bootup()                                           # nothing recorded yet: warns, boots
    assert seen == []
    wrote = node.boot_self_test(write=True)                 # `cli golden --write`
    key = environment_key("kandinsky2")
    assert wrote == {key: solution_cid({"out-1.png": PNG})} and store.get(key) == wrote[key]
    assert seen[-1]["prompt"] == "arbius test cat" and seen[-1]["seed"] == 1337
    node.bootup()                                           # matches
    store.put(key, "0x1220" + "ab" * 32)                    # a different build / a mis-built library
    with pytest.raises(FatalMinerError) as ei:
        node.bootup()
    assert ei.value.code == 1                               # process.exit(1) semantics
    # another GPU model or library version is another key: not silently compared against this golden
    assert environment_key("anythingv3") != key


def test_cli_golden_write_then_check(tmp_path, capsys):
    from arbius_b200 import cli
    cfgp = tmp_path / "MiningConfig.json"
    cfgp.write_text(json.dumps({"db_path": str(tmp_path / "db.sqlite"), "blockchain": {"rpc_url": "sim://"},
                                "rpc": {"host": "127.0.0.1", "port": 0},
                                "ml": {"strategy": "cog", "cog": {}}, "ipfs": {"strategy": "local"},
                                "selftest": {"goldens": str(tmp_path / "g.json")}}))
    assert cli.main(["golden", str(cfgp)]) == 0
    assert "no goldens recorded" in capsys.readouterr().out


@pytest.mark.parametrize("payload", [b'{"prompt": "p", "x": NaN}', b'{"prompt": "p", "x": -Infinity}', b"null"])
def test_inputs_json_parse_rejects_are_invalid_tasks(tmp_path, payload):
    calls = []
    eng, node, mid = _node(tmp_path, lambda inp: (calls.append(inp), PNG)[1])
    node.bootup()
    tid = eng.submit_task(USER, 0, USER, mid, 0, payload)
    node.run_until_idle()
    assert calls == [] and node.db.get_invalid_task(tid) is not None and tid not in eng.solutions


def test_failed_solve_is_replayed_a_bounded_number_of_times(tmp_path):
    n = {"calls": 0}

    def flaky(inp):
        n["calls"] += 1
        if n["calls"] <= 1:
            raise RuntimeError("worker restarted")
        return PNG
    eng, node, mid = _node(tmp_path, flaky, solve_retries=2)
    node._retry = lambda fn, tries=10, base=1.5: fn()        # no inner retries: the job itself must come back
    node.bootup()
    tid = eng.submit_task(USER, 0, USER, mid, 0, b'{"prompt": "p"}')
    node.run_until_idle()
    eng.time_travel(30)
    node.run_until_idle()
    assert tid in eng.solutions and node.db.get_failed_jobs() == []

    def dead(inp):
        raise RuntimeError("gone")
    eng2, node2, mid2 = _node(tmp_path, dead, solve_retries=1)
    node2._retry = lambda fn, tries=10, base=1.5: fn()
    node2.bootup()
    eng2.submit_task(USER, 0, USER, mid2, 0, b'{"prompt": "p"}')
    for _ in range(4):
        node2.run_until_idle()
        eng2.time_travel(30)
    failed = node2.db.get_failed_jobs()
    assert len(failed) == 1 and json.loads(failed[0]["data"])["_retries"] == 1


def test_concurrent_sends_from_one_key_never_share_a_nonce():
    from arbius_b200.chain import contract_abi as CA
    from arbius_b200.chain import jsonrpc

    class FakeProvider:
        def __init__(self):
            self.used, self.lock, self.mined = [], threading.Lock(), 0

        def chain_id(self):
            return 31337

        def nonce(self, addr):
            time.sleep(0.002)                               # widen the read-then-send window
            return self.mined

        def gas_price(self):
            return 1

        def request(self, method, params):
            assert method == "eth_sendRawTransaction"
            from arbius_b200.chain.eth import decode_raw_transaction as decode_transaction
            tx = decode_transaction(bytes.fromhex(params[0][2:]))
            with self.lock:
                if tx.nonce in self.used:
                    raise jsonrpc.JsonRpcError(-32000, "nonce too low")
                self.used.append(tx.nonce)
                self.mined = max(self.mined, tx.nonce + 1)
            return "0x" + "00" * 32
    p = FakeProvider()
    key = "0x" + "11" * 32
    a = jsonrpc.Contract(p, "0x" + "aa" * 20, CA.ENGINE_FUNCTIONS, CA.ENGINE_EVENTS, key, 31337)
    b = jsonrpc.Contract(p, "0x" + "bb" * 20, CA.TOKEN_FUNCTIONS, CA.TOKEN_EVENTS, key, 31337)
    errs = []

    def worker(c, name, *args):
        try:
            for _ in range(10):
                c.send(name, *args, wait=False)
        except Exception as e:  # noqa: BLE001
            errs.append(e)
    ts = [threading.Thread(target=worker, args=(a, "signalCommitment", b"\x01" * 32)) for _ in range(3)]
    ts += [threading.Thread(target=worker, args=(b, "approve", "0x" + "cc" * 20, 1))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert errs == [] and sorted(p.used) == list(range(40))


def test_batch_scale_is_per_thread():
    from fractions import Fraction
    import arbius_b200.ops  # noqa: F401
    G = sys.modules["arbius_b200.ops.gemm"]
    seen = {}
    go, done = threading.Event(), threading.Event()

    def other():
        go.wait(5)
        seen["other"] = G.get_batch_scale()                 # the capture thread: must still see 1
        done.set()
    t = threading.Thread(target=other)
    t.start()
    with G.batch_scale(Fraction(1, 3)):
        go.set()
        done.wait(5)
        assert G._canon(6) == 2
        with G.batch_scale(2):
            assert G._canon(1) == 2
        assert G.get_batch_scale() == Fraction(1, 3)
    t.join()
    assert seen["other"] == 1 and G.get_batch_scale() == 1


def test_worker_treats_device_faults_as_fatal_and_plain_errors_as_422():
    import urllib.error
    import urllib.request
    from arbius_b200.worker import server as ws
    assert ws.is_fatal_device_error(RuntimeError("CUDA error: an illegal memory access was encountered"))
    assert not ws.is_fatal_device_error(ValueError("width must be one of ..."))
    fatal = []
    mode = {"m": "bad"}

    def predict(inp):
        if mode["m"] == "bad":
            raise ValueError("bad input")
        raise RuntimeError("CUDA error: unspecified launch failure")
    srv = ws.make_server(predict, load_template("kandinsky2"), "127.0.0.1", 0, on_fatal=fatal.append)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    url = f"http://127.0.0.1:{srv.server_address[1]}/predictions"

    def post():
        req = urllib.request.Request(url, json.dumps({"input": {"prompt": "p", "seed": 1}}).encode(),
                                     {"Content-Type": "application/json"})
        try:
            return urllib.request.urlopen(req, timeout=5).status
        except urllib.error.HTTPError as e:
            return e.code
    assert post() == 422 and fatal == []
    mode["m"] = "cuda"
    assert post() == 500
    for _ in range(100):                                     # the handler answers first, then reports the fault
        if fatal:
            break
        time.sleep(0.02)
    assert len(fatal) == 1
    srv.shutdown()
