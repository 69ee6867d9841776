"""Miner control plane against EngineSim with a fake worker: BASELINE config 1 (Engine task lifecycle,
submitTask -> signalCommitment -> submitSolution -> claimSolution with CIDv0 of a fixed output, CPU only),
plus the pieces the reference never tested: hydration failures, contestation, scheduler, RPC, worker."""
import json
import threading
import time
import urllib.request

import pytest

from arbius_b200.chain.client import SimChainClient
from arbius_b200.chain.engine_sim import UNIT, ZERO_ADDR, deploy_local
from arbius_b200.miner import rpc as miner_rpc
from arbius_b200.miner.db import MinerDB
from arbius_b200.miner.ml import (CogBackend, MLError, NativeBackend, ReplicateBackend, decode_cog_output,
                                  encode_cog_output)
from arbius_b200.miner.node import EVIL_CID, FatalMinerError, MinerNode
from arbius_b200.miner.scheduler import CostModel, GpuPool, SubsetPlanner, simulate
from arbius_b200.miner.utils import expretry
from arbius_b200.protocol import (MiningFilter, ModelEntry, load_template, solution_cid, taskid2seed, template_bytes)
from arbius_b200.protocol.config import parse_mining_config
from arbius_b200.worker import server as worker_server

MINER, USER, OTHER = "0x" + "a1" * 20, "0x" + "11" * 20, "0x" + "a2" * 20
FIXED_PNG = b"\x89PNG\r\n\x1a\n" + b"fixed-output" * 10


def _setup(evil=False, n_gpus=1, predictor=None, supply="599990"):
    eng = deploy_local(engine_balance=int(supply) * UNIT, funded={MINER: 50 * UNIT, USER: 50 * UNIT, OTHER: 50 * UNIT})
    mid = eng.register_model(USER, USER, 0, template_bytes("anythingv3"))
    model = ModelEntry(id=mid, template=load_template("anythingv3"), filters=[MiningFilter()], name="anythingv3")
    calls = []

    def predict(inp):
        calls.append(inp)
        return FIXED_PNG if predictor is None else predictor(inp)
    cfg = parse_mining_config({"db_path": ":memory:", "evilmode": evil, "stake_buffer_percent": 20,
                               "stake_buffer_topup_percent": 1, "ml": {"strategy": "native"},
                               "ipfs": {"strategy": "local"}, "automine": {"enabled": False}})
    ml = NativeBackend({mid: [predict] * n_gpus})
    chain = SimChainClient(eng, MINER)
    node = MinerNode(cfg, chain, ml, [model], db=MinerDB(":memory:"), n_gpus=n_gpus, sleep=lambda s: None)
    node.bootup()
    return eng, node, mid, calls


def _submit(eng, mid, inp, fee=0, sender=USER):
    eng.token.approve(sender, eng.address, fee)
    return eng.submit_task(sender, 0, sender, mid, fee, json.dumps(inp, separators=(",", ":")).encode())


def test_full_lifecycle_commit_reveal_claim():
    eng, node, mid, calls = _setup(supply="597000")         # supply 3000 -> validators must stake 2.4
    node.run_until_idle()                                    # validatorStake job tops the stake up
    assert eng.validator(MINER).staked == 3 * UNIT           # min 2.4 * 100/80
    tid = _submit(eng, mid, {"prompt": "arbius test cat", "negative_prompt": ""}, fee=UNIT)
    node.run_until_idle()
    # the model saw the hydrated input with defaults filled and the seed derived from the task id
    assert len(calls) == 1
    assert calls[0]["width"] == 768 and calls[0]["scheduler"] == "DPMSolverMultistep"
    assert calls[0]["seed"] == taskid2seed(tid) and list(calls[0])[-1] == "seed"
    sol = eng.solutions[tid]
    assert sol.validator == MINER and sol.cid == solution_cid({"out-1.png": FIXED_PNG})
    names = [e.name for e in eng.events]
    assert names.index("SignalCommitment") < names.index("SolutionSubmitted")
    # claim job is delayed by 2000 + 120 s
    claim = [j for j in node.db.get_jobs() if j.method == "claim"]
    assert len(claim) == 1 and claim[0].waituntil >= sol.blocktime + 2120 - 2 and claim[0].priority == 50
    before = eng.token.balance_of(MINER)
    eng.time_travel(2200)
    node.run_until_idle()
    assert eng.solutions[tid].claimed and eng.token.balance_of(MINER) == before + UNIT * 9 // 10
    assert node.db.get_failed_jobs() == []
    assert node.solved[0]["taskid"] == tid and node.solved[0]["total_ms"] >= 0


def test_same_input_same_cid_twice_and_evilmode():
    eng, node, mid, _ = _setup()
    a = _submit(eng, mid, {"prompt": "x", "negative_prompt": "y"})
    b = _submit(eng, mid, {"prompt": "x", "negative_prompt": "y"})
    node.run_until_idle()
    assert eng.solutions[a].cid == eng.solutions[b].cid
    eng2, node2, mid2, calls2 = _setup(evil=True)
    t = _submit(eng2, mid2, {"prompt": "x", "negative_prompt": "y"})
    node2.run_until_idle()
    assert eng2.solutions[t].cid == EVIL_CID and calls2 == []


def test_invalid_inputs_are_recorded_and_not_solved():
    eng, node, mid, calls = _setup()
    bad_json = eng.submit_task(USER, 0, USER, mid, 0, b"\xff\xfenot json")
    missing = _submit(eng, mid, {"prompt": "only prompt"})
    decimal = _submit(eng, mid, {"prompt": "p", "negative_prompt": "", "guidance_scale": 7.5})
    version1 = eng.submit_task(USER, 1, USER, mid, 0, b"{}")
    node.run_until_idle()
    assert calls == []
    for t in (bad_json, missing, decimal, version1):
        assert node.db.get_invalid_task(t) is not None and t not in eng.solutions


def test_contests_wrong_solution_to_invalid_task_and_losing_race():
    eng, node, mid, _ = _setup(supply="597000")
    node.run_until_idle()
    for a in (OTHER,):
        eng.token.approve(a, eng.address, 3 * UNIT)
        eng.validator_deposit(a, a, 3 * UNIT)
    # an invalid task that somebody else "solves": we must contest (index.ts:253-256)
    bad = _submit(eng, mid, {"prompt": "only prompt"})
    node.run_until_idle()
    from arbius_b200.protocol import generate_commitment
    eng.signal_commitment(OTHER, generate_commitment(OTHER, bad, EVIL_CID))
    eng.submit_solution(OTHER, bad, EVIL_CID)
    node.run_until_idle()
    assert eng.contestations[bad].validator == MINER
    assert any(j.method == "contestationVoteFinish" for j in node.db.get_jobs())
    # losing the race with a different CID -> contest as well (index.ts:663-669)
    good = _submit(eng, mid, {"prompt": "p", "negative_prompt": ""})
    node.process_task(good, eng.task_txs[good])
    eng.signal_commitment(OTHER, generate_commitment(OTHER, good, EVIL_CID))
    eng.submit_solution(OTHER, good, EVIL_CID)
    node.db.clear_jobs_by_method("solve")
    # pretend we had not seen the solution yet when we started solving
    orig = node.chain.solutions
    state = {"n": 0}

    def first_call_empty(t):
        state["n"] += 1
        from arbius_b200.chain.client import SolutionView
        return SolutionView(ZERO_ADDR, 0, False, "0x") if state["n"] == 1 else orig(t)
    node.chain.solutions = first_call_empty
    node.process_solve(good)
    node.chain.solutions = orig
    assert eng.contestations[good].validator == MINER
    # after the vote period the finish job releases stakes
    eng.time_travel(5100)
    node.run_until_idle()
    assert eng.contestations[bad].finish_start_index > 0


def test_version_mismatch_is_fatal():
    eng, node, mid, _ = _setup()
    eng.set_version(eng.owner, 1)
    with pytest.raises(FatalMinerError):
        node.step()                               # VersionChanged event -> version_check -> exit


def test_concurrent_solves_use_all_gpus():
    gate = threading.Barrier(4, timeout=10)

    def slow(inp):
        gate.wait()                                # only passes if 4 predictions are in flight at once
        return FIXED_PNG + str(inp["seed"]).encode()
    eng, node, mid, calls = _setup(n_gpus=4, predictor=slow)
    tids = [_submit(eng, mid, {"prompt": f"p{i}", "negative_prompt": ""}) for i in range(4)]
    node.run_until_idle()
    assert len(calls) == 4 and all(t in eng.solutions for t in tids)
    assert len({eng.solutions[t].cid for t in tids}) == 4
    assert sorted(s["gpus"] for s in node.solved) == [(0,), (1,), (2,), (3,)]


def test_jobs_survive_failure_and_can_be_replayed():
    def boom(inp):
        raise RuntimeError("gpu fell off the bus")
    eng, node, mid, _ = _setup(predictor=boom)
    node._retry = lambda fn, tries=10, base=1.5: expretry(fn, 1, 1.0, sleep=lambda s: None)
    tid = _submit(eng, mid, {"prompt": "p", "negative_prompt": ""})
    node.run_until_idle()
    assert tid not in eng.solutions
    db = node.db
    db.store_failed_job(db.queue_job("solve", 20, 0, False, {"taskid": tid}))
    assert db.requeue_failed_jobs() >= 1 and db.get_failed_jobs() == []


# ---------------------------------------------------------------- scheduler
def test_subset_planner_alignment_and_policies():
    p = SubsetPlanner(8)
    assert p.acquire(4) == (0, 1, 2, 3) and p.acquire(2) == (4, 5) and p.acquire(4) is None
    assert p.largest_free_aligned() == 2 and p.acquire(1) == (6,)
    p.release((0, 1, 2, 3))
    assert p.largest_free_aligned() == 4
    cost = CostModel({"sd": {1: 136.0, 2: 80.0, 4: 50.0, 8: 35.0}, "video": {1: 9000.0, 2: 4800.0, 4: 2600.0, 8: 1500.0}},
                     efficiency=0.8)
    assert CostModel().best_latency_p("unmeasured", 8) == 1          # never widen without evidence
    q = SubsetPlanner(8)
    assert q.choose_p("sd", waiting=20, cost=cost, policy="throughput") == 1     # backlog: one GPU per task
    assert q.choose_p("video", waiting=1, cost=cost, policy="throughput") == 8   # lone task: whole box
    assert q.choose_p("video", waiting=2, cost=cost, policy="latency") == 4
    assert cost.best_throughput_p("sd", 8) == 1 and cost.ms("sd", 16) < 35.0


def test_scheduler_simulation_beats_serial_reference():
    cost = CostModel({"sd": {1: 136.0, 2: 80.0, 4: 50.0, 8: 35.0}})
    burst = [(0.0, "sd")] * 64
    serial = simulate(burst, 1, cost, "throughput")
    packed = simulate(burst, 8, cost, "throughput")
    assert packed["tasks_per_hour"] > 7.5 * serial["tasks_per_hour"]
    trickle = [(i * 1.0, "sd") for i in range(20)]         # one task per second: latency mode widens tasks
    lat = simulate(trickle, 8, cost, "latency")
    thr1 = simulate(trickle, 1, cost, "throughput")
    assert lat["p50_latency_s"] < 0.5 * thr1["p50_latency_s"]


def test_scheduler_subset_reaches_the_backend_and_pairs_run_on_both_gpus():
    """A lone task under the latency policy gets an aligned GPU pair when a pair predictor with a measured gain exists;
    the backend then runs THAT subset's predictor while holding both GPUs; with a full queue tasks go one per GPU."""
    eng = deploy_local(engine_balance=599990 * UNIT, funded={MINER: 50 * UNIT, USER: 50 * UNIT})
    mid = eng.register_model(USER, USER, 0, template_bytes("anythingv3"))
    model = ModelEntry(id=mid, template=load_template("anythingv3"), filters=[MiningFilter()], name="anythingv3")
    single_calls, pair_calls = [], []
    ml = None

    def single(g):
        def f(inp):
            single_calls.append(g)
            return FIXED_PNG + str(inp["seed"]).encode()
        return f

    def pair(inp):
        pair_calls.append([lk.locked() for lk in ml._locks[mid.lower()]])
        return FIXED_PNG + str(inp["seed"]).encode()
    ml = NativeBackend({mid: [single(0), single(1)]}, {mid: {(0, 1): pair}})
    assert ml.subset_sizes(model) == [1, 2]
    cfg = parse_mining_config({"db_path": ":memory:", "ml": {"strategy": "native"}, "ipfs": {"strategy": "local"},
                               "automine": {"enabled": False}, "scheduler": {"policy": "latency"}})
    node = MinerNode(cfg, SimChainClient(eng, MINER), ml, [model], db=MinerDB(":memory:"), n_gpus=2,
                     cost=CostModel.measured_b200(["anythingv3"]), sleep=lambda s: None)
    node.bootup()
    node.run_until_idle()
    tid = _submit(eng, mid, {"prompt": "alone", "negative_prompt": ""})
    node.run_until_idle()
    assert pair_calls == [[True, True]] and single_calls == []
    assert node.solved[-1]["gpus"] == (0, 1) and tid in eng.solutions
    # explicit single-GPU subsets go to that GPU's predictor
    ml.getfiles(model, "0x0", {"seed": 1}, gpus=(1,))
    assert single_calls == [1]
    # without a measured pair row nothing is widened
    assert CostModel.measured_b200([]).best_latency_p("anythingv3", 2) == 1
    assert CostModel.measured_b200(["zeroscopev2xl"]).best_latency_p("zeroscopev2xl", 8) == 2


def test_gpu_pool_runs_and_learns_costs():
    pool = GpuPool(2, CostModel(), policy="throughput")
    futs = [pool.submit("sd", lambda gpus, i=i: (time.sleep(0.01), i, gpus)[1:], priority=1) for i in range(6)]
    assert pool.drain(10)
    assert sorted(f.result()[0] for f in futs) == list(range(6))
    assert "sd" in pool.cost.table and len(pool.completed) == 6


# ---------------------------------------------------------------- adapters
def test_cog_contract_round_trip_and_backends():
    t = load_template("anythingv3")
    res = encode_cog_output(FIXED_PNG, "image/png")
    assert res["output"][0].startswith("data:image/png;base64,") and decode_cog_output(res) == FIXED_PNG
    with pytest.raises(MLError):
        decode_cog_output({"output": []})
    m = ModelEntry(id="0xAB", template=t)
    seen = []
    cog = CogBackend({"0xab": {"url": "http://w0", "urls": ["http://w0", "http://w1"]}},
                     post=lambda url, payload, timeout: (seen.append((url, payload)), res)[1])
    assert cog.endpoints(m) == 2
    assert cog.getfiles(m, "0x1", {"prompt": "p", "seed": 1}) == {"out-1.png": FIXED_PNG}
    cog.getfiles(m, "0x2", {"prompt": "p", "seed": 2})
    assert [u for u, _ in seen] == ["http://w0", "http://w1"] and seen[0][1] == {"input": {"prompt": "p", "seed": 1}}
    assert ReplicateBackend.model_ref(t).startswith("cjwbw/anything-v3-better-vae:09a5805")
    with pytest.raises(MLError):
        ReplicateBackend(None).getfiles(m, "0x1", {})


def test_worker_server_speaks_cog_contract():
    t = load_template("anythingv3")
    srv = worker_server.make_server(lambda inp: FIXED_PNG + str(inp["seed"]).encode(), t, "127.0.0.1", 0)
    port = srv.server_address[1]
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        req = urllib.request.Request(f"http://127.0.0.1:{port}/predictions", data=json.dumps({"input": {"prompt": "p", "seed": 7}}).encode(),
                                     headers={"Content-Type": "application/json"})
        body = json.loads(urllib.request.urlopen(req, timeout=5).read())
        assert decode_cog_output(body) == FIXED_PNG + b"7"
        cog = CogBackend({"0xab": {"url": f"http://127.0.0.1:{port}/predictions"}})
        assert cog.getfiles(ModelEntry(id="0xab", template=t), "0x1", {"prompt": "p", "seed": 7}) == {"out-1.png": FIXED_PNG + b"7"}
        bad = urllib.request.Request(f"http://127.0.0.1:{port}/predictions", data=b'{"input": {"prompt": "p"}}')
        with pytest.raises(urllib.error.HTTPError):
            urllib.request.urlopen(bad, timeout=5)
    finally:
        srv.shutdown()


def test_operator_rpc_routes():
    db = MinerDB(":memory:")
    st, body, _ = miner_rpc.handle(db, "GET", "/", {})
    assert st == 200 and body == "Arbius Miner RPC"
    ok = {"method": "solve", "priority": 20, "waituntil": 0, "concurrent": False, "data": {"taskid": "0x1"}}
    assert miner_rpc.handle(db, "POST", "/api/jobs/queue", ok)[1] == {"status": "ok"}
    bad = dict(ok, priority="20")
    assert miner_rpc.handle(db, "POST", "/api/jobs/queue", bad)[1]["status"] == "fail"
    assert "missing required field (data)" in miner_rpc.handle(db, "POST", "/api/jobs/queue", {k: v for k, v in ok.items() if k != "data"})[1]["e"]
    jobs = miner_rpc.handle(db, "POST", "/api/jobs/get", {"limit": 10})[1]["jobs"]
    assert len(jobs) == 1 and jobs[0]["method"] == "solve"
    assert miner_rpc.handle(db, "POST", "/api/jobs/list", {})[1]["status"] == "ok"
    assert miner_rpc.handle(db, "POST", "/api/jobs/delete", {"id": jobs[0]["id"]})[1] == {"status": "ok"}
    assert miner_rpc.handle(db, "POST", "/api/jobs/delete", {"id": "x"})[1]["status"] == "fail"
    assert db.get_jobs() == []
    assert "arbius_jobs_queued 0" in miner_rpc.handle(db, "GET", "/metrics", {})[1]


def test_db_job_order_and_seed_reinjection():
    db = MinerDB(":memory:")
    db.queue_job("automine", 5)
    db.queue_job("claim", 50, waituntil=99)
    db.queue_job("solve", 20)
    assert [j.method for j in db.get_jobs()] == ["claim", "solve", "automine"]
    tid = "0x" + "ab" * 32
    db.store_task_input(tid, "0xcid", {"prompt": "p"})
    assert db.get_task_input(tid, "0xcid") == {"prompt": "p", "seed": taskid2seed(tid)}
    db.clear_jobs_by_method("solve")
    assert [j.method for j in db.get_jobs()] == ["claim", "automine"]


def test_expretry_returns_none_after_exhaustion():
    n = {"c": 0}

    def flaky():
        n["c"] += 1
        if n["c"] < 3:
            raise IOError("nope")
        return 42
    waits = []
    assert expretry(flaky, sleep=waits.append) == 42 and waits == [1.0, 1.5]
    assert expretry(lambda: 1 / 0, tries=2, sleep=lambda s: None) is None


def test_task_slots_let_a_batching_predictor_merge_solves():
    """slots=2: two solves are admitted per GPU and the batching predictor runs them as one group; subsets of two GPUs
    still need both GPUs idle."""
    from arbius_b200.models.registry import BatchingPredictor
    pl = SubsetPlanner(2, slots=2)
    assert [pl.acquire(1) for _ in range(5)] == [(0,), (1,), (0,), (1,), None]
    assert pl.largest_free_aligned() == 0 and pl.free_count() == 0
    pl.release((0,)); pl.release((0,))
    assert pl.largest_free_aligned() == 1 and pl.acquire(2) is None          # GPU 1 still has work
    pl.release((1,)); pl.release((1,))
    assert pl.acquire(2) == (0, 1) and pl.acquire(1) is None
    pl.release((0, 1))
    assert pl.free_count() == 4

    class FakePipe:
        groups = []

        @staticmethod
        def _task_args(inp):
            return dict(width=64, height=64, num_inference_steps=int(inp.get("num_inference_steps", 2)),
                        guidance_scale=7.0, scheduler="x", seed=inp["seed"])

        def generate_batch(self, tasks):
            self.groups.append([t["seed"] for t in tasks])
            time.sleep(0.02)
            return [bytes([t["seed"]]) for t in tasks]
    bp = BatchingPredictor(FakePipe(), None, task_batch=2, window_ms=200.0, encode=lambda img: FIXED_PNG + img)
    eng, node, mid, _ = _setup()
    model = node.models[0]
    ml = NativeBackend({mid: [bp]}, slots=2)
    out = {}
    ths = [threading.Thread(target=lambda s=s: out.__setitem__(s, ml.getfiles(model, "0x0", {"seed": s}, gpus=(0,))))
           for s in (1, 2, 3)]
    for t in ths:
        t.start()
    for t in ths:
        t.join(10)
    assert {s: v["out-1.png"] for s, v in out.items()} == {s: FIXED_PNG + bytes([s]) for s in (1, 2, 3)}
    assert sorted(len(g) for g in FakePipe.groups) == [1, 2]                    # two merged, the third ran after them
    # a request with other settings is not merged into the group
    FakePipe.groups.clear()
    ths = [threading.Thread(target=lambda kw=kw: ml.getfiles(model, "0x0", kw, gpus=(0,)))
           for kw in ({"seed": 7}, {"seed": 8, "num_inference_steps": 5})]
    for t in ths:
        t.start()
    for t in ths:
        t.join(10)
    assert sorted(FakePipe.groups) == [[7], [8]]
    bp.close()


def test_cost_table_widens_video_tasks_over_idle_gpus():
    cost = CostModel.measured_b200([], ["zeroscopev2xl"], [2, 4, 8])
    assert cost.best_latency_p("zeroscopev2xl", 8) == 8 and cost.best_latency_p("zeroscopev2xl", 4) == 4
    assert cost.best_throughput_p("zeroscopev2xl", 8) == 1               # 8 GPUs on one clip never beat 8 clips at once
    pl = SubsetPlanner(8)
    assert pl.choose_p("zeroscopev2xl", 1, cost, "latency") == 8          # a lone clip takes the box
    assert pl.choose_p("zeroscopev2xl", 2, cost, "latency") == 4          # two clips: half each
    assert pl.choose_p("zeroscopev2xl", 8, cost, "throughput") == 1
    # sizes that were not built stay out of the table
    assert CostModel.measured_b200([], ["zeroscopev2xl"], [2]).best_latency_p("zeroscopev2xl", 8) == 2
    sim = simulate([(0.0, "zeroscopev2xl")], 8, cost, "latency")
    assert abs(sim["p50_latency_s"] - CostModel.MEASURED_B200_FRAMES["zeroscopev2xl"][8] / 1e3) < 1e-6


def test_subset_predictor_out_of_heap_falls_back_to_one_gpu():
    eng, node, mid, _ = _setup()
    model = node.models[0]

    def too_big(inp):
        raise MemoryError("peer heap exhausted")
    ml = NativeBackend({mid: [lambda inp: FIXED_PNG + b"0", lambda inp: FIXED_PNG + b"1"]}, {mid: {(0, 1): too_big}})
    assert ml.getfiles(model, "0x0", {"seed": 1}, gpus=(0, 1)) == {"out-1.png": FIXED_PNG + b"0"}
    assert not any(lk.locked() for lk in ml._locks[mid.lower()])


def test_subset_predictor_unsupported_shape_falls_back_but_other_errors_propagate():
    from arbius_b200.models.video.pipeline import UnsupportedSplit
    eng, node, mid, _ = _setup()
    model = node.models[0]

    def no_split(inp):
        raise UnsupportedSplit("10 frames over 4 ranks")

    def broken(inp):
        raise ValueError("bad weights")
    ml = NativeBackend({mid: [lambda inp: FIXED_PNG, lambda inp: FIXED_PNG]}, {mid: {(0, 1): no_split}})
    assert ml.getfiles(model, "0x0", {"seed": 1}, gpus=(0, 1)) == {"out-1.png": FIXED_PNG}
    ml = NativeBackend({mid: [lambda inp: FIXED_PNG, lambda inp: FIXED_PNG]}, {mid: {(0, 1): broken}})
    with pytest.raises(ValueError, match="bad weights"):
        ml.getfiles(model, "0x0", {"seed": 1}, gpus=(0, 1))
    assert not any(lk.locked() for lk in ml._locks[mid.lower()])
