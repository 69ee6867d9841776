"""MinerNode — event handlers, job processors and the main loop of the mining control plane.

Mirrors the reference flow (reference: miner/src/index.ts; call stacks in SURVEY §3.1-3.5):
chain events -> durable jobs -> ``task`` (filter, hydrate input, derive seed) -> ``solve`` (run model,
hash outputs to a CIDv0, signalCommitment, submitSolution) -> delayed ``claim``; plus the periodic
``validatorStake`` / ``automine`` jobs, contestation handling and the boot self-test. Differences:
  * ``solve`` jobs run concurrently on the GPU scheduler (one per GPU subset), not serially;
  * a job row is deleted only after the job has finished (a crash re-runs it instead of losing it);
  * outputs are per-task byte strings, never a shared ``cache/out-1.png``;
  * time comes from the chain client so the whole node can be driven deterministically in tests.
Job methods, priorities and payloads are unchanged.
"""
from __future__ import annotations

import json
import threading
import time
from concurrent.futures import Future
from typing import Callable, Dict, List, Optional, Sequence

from ..chain.client import ChainClient, ChainError
from ..chain.engine_sim import ZERO_ADDR, Event
from ..protocol.abi import generate_commitment
from ..protocol.base58 import b58decode
from ..protocol.config import MiningConfig
from ..protocol.hydrate import ModelEntry, check_model_filter, get_model_by_id, hydrate_input
from ..protocol.seed import taskid2seed
from . import ipfs
from .db import Job, MinerDB
from .ml import MLBackend
from .scheduler import CostModel, GpuPool
from .utils import expretry, log

EVIL_CID = "0x1220" + "6" * 64          # reference: miner/src/models.ts:40-42
CLAIM_DELAY = 2000 + 120                 # index.ts:644
CONTESTATION_FINISH_DELAY = 4000 + 1010  # index.ts:683
MAX_UINT256 = (1 << 256) - 1


class FatalMinerError(SystemExit):
    """Conditions on which the reference calls process.exit(1)."""


def _reject_json_constant(name: str):
    raise ValueError(f"{name} is not valid JSON")


class MinerNode:
    def __init__(self, config: MiningConfig, chain: ChainClient, ml: MLBackend, models: Sequence[ModelEntry],
                 db: Optional[MinerDB] = None, n_gpus: int = 1, cost: Optional[CostModel] = None,
                 sleep: Callable[[float], None] = time.sleep, retry_sleep: Optional[Callable[[float], None]] = None,
                 pin: bool = False, slots: int = 1, goldens=None, solve_retries: int = 2):
        self.c = config
        self.chain = chain
        self.ml = ml
        self.models = list(models)
        self.db = db or MinerDB(config.db_path)
        self.pool = GpuPool(n_gpus, cost, policy=config.scheduler.policy, slots=slots)   # slots: solves per GPU at once
        self._sleep = sleep
        self._retry_sleep = retry_sleep or sleep
        self.pin = pin
        self.goldens = goldens                  # GoldenStore: when set, bootup() runs the determinism self-test
        self.solve_retries = solve_retries      # automatic replays of a failed solve before it is dead-lettered
        self._inflight: Dict[int, Future] = {}
        self._lock = threading.RLock()
        self.solved: List[dict] = []            # per-task latency log (metrics)
        self.stopped = False

    # ------------------------------------------------------------------ helpers
    def _retry(self, fn, tries: int = 10, base: float = 1.5):
        return expretry(fn, tries, base, sleep=self._retry_sleep)

    def lookup_and_insert_task(self, taskid: str):
        existing = self.db.get_task(taskid)
        if existing:
            from ..chain.client import TaskView
            return TaskView(existing["modelid"], int(existing["fee"]), existing["address"], int(existing["blocktime"]),
                            existing["version"], existing["cid"])
        t = self._retry(lambda: self.chain.tasks(taskid))
        if t is None:
            return None
        self.db.store_task(taskid, t.model, t.fee, t.owner, t.blocktime, t.version, t.cid)
        return t

    def lookup_and_insert_task_input(self, taskid: str, cid: str, txid: str, template: dict) -> bool:
        """index.ts:131-189: decode calldata -> JSON -> hydrate -> seed -> store (+ pin job)."""
        if self.db.get_task_input(taskid, cid) is not None:
            return True
        raw = self._retry(lambda: self.chain.task_input_bytes(taskid, txid))
        if raw is None:
            return False
        try:
            # JSON.parse semantics (index.ts:160-164): NaN / Infinity are syntax errors, and indexing a non-object top
            # level throws inside hydrateInput — both make the task invalid for every reference miner, so here too
            pre = json.loads(raw.decode("utf-8"), parse_constant=_reject_json_constant)
            if pre is None:
                raise ValueError("null input")
        except (UnicodeDecodeError, ValueError):
            log.warning("Task (%s) unable to parse input", taskid)
            self.db.store_invalid_task(taskid)
            return False
        hyd = hydrate_input(pre, template)
        if hyd.err:
            log.warning("Task (%s) hydration error %s", taskid, hyd.errmsg)
            self.db.store_invalid_task(taskid)
            return False
        inp = dict(hyd.input)
        inp["seed"] = taskid2seed(taskid)
        self.db.store_task_input(taskid, cid, inp)
        self.db.queue_job("pinTaskInput", priority=10, waituntil=0, concurrent=True,
                          data={"taskid": taskid, "input": json.dumps(pre)})
        return True

    # ------------------------------------------------------------------ event handlers (index.ts:191-333)
    def on_event(self, ev: Event) -> None:
        h = getattr(self, "_on_" + ev.name, None)
        if h:
            h(ev)

    def _on_VersionChanged(self, ev):
        self.version_check()

    def _on_TaskSubmitted(self, ev):
        taskid = ev.args["id"]
        self.lookup_and_insert_task(taskid)
        self.db.queue_job("task", priority=10, waituntil=0, concurrent=True, data={"taskid": taskid, "txid": ev.txhash})

    def _on_TaskRetracted(self, ev):
        taskid = ev.args["id"]
        if not self.db.get_task(taskid):
            self.lookup_and_insert_task(taskid)
        self.db.update_task_set_retracted(taskid)

    def _on_SolutionSubmitted(self, ev):
        taskid = ev.args["task"]
        if self.db.get_solution(taskid):
            return
        s = self._retry(lambda: self.chain.solutions(taskid))
        if s is None:
            return
        if self.db.get_invalid_task(taskid) is not None and s.validator != self.chain.address:
            self.contest_solution(taskid)
        self.db.store_solution(taskid, s.validator, s.blocktime, s.claimed, s.cid)

    def _on_ContestationSubmitted(self, ev):
        taskid = ev.args["task"]
        if self.db.get_contestation(taskid):
            return
        c = self._retry(lambda: self.chain.contestations(taskid))
        if c is None:
            return
        if self.db.get_invalid_task(taskid) is not None:
            self.vote_on_contestation(taskid, True)
        self.db.store_contestation(taskid, c.validator, c.blocktime, c.finish_start_index)

    def _on_ContestationVote(self, ev):
        self.db.store_contestation_vote(ev.args["task"], ev.args["addr"], bool(ev.args["yea"]))

    # ------------------------------------------------------------------ job processors
    def process_pin_task_input(self, taskid: str, inp: str) -> None:
        """index.ts:379-390."""
        cid = self._retry(lambda: ipfs.pin_file(self.c, inp.encode(), "input", pin=self.pin))
        log.debug("Task input %s pinned with %s", taskid, cid)

    def process_validator_stake(self) -> None:
        """index.ts:397-472."""
        if self.chain.eth_balance() < 10 ** 16:
            log.warning("BCHK Low Ether balance")
        staked = self.chain.validator_staked()
        vmin = self._retry(self.chain.get_validator_minimum)
        self.db.queue_job("validatorStake", priority=100, waituntil=self.chain.now() + 600, concurrent=False,
                          data={"validatorMinimum": str(vmin)})
        min_topup = vmin * 100 // (100 - self.c.stake_buffer_topup_percent)
        if staked >= min_topup:
            log.debug("BCHK Have sufficient stake")
            return
        deposit = vmin * 100 // (100 - self.c.stake_buffer_percent) - staked
        balance = self._retry(self.chain.token_balance)
        if balance < deposit:
            log.error("BCHK Balance %d less than deposit amount %d", balance, deposit)
            raise RuntimeError("unable to stake required balance")
        allowance = self._retry(lambda: self.chain.token_allowance(self.chain.address, self.chain.solver_address))
        if allowance < balance:
            self._retry(lambda: self.chain.token_approve(self.chain.solver_address, MAX_UINT256 - allowance))
        self._retry(lambda: self.chain.validator_deposit(self.chain.address, deposit))
        log.info("BCHK Post staked: %d", self.chain.validator_staked())

    def process_automine(self) -> None:
        """index.ts:474-503."""
        a = self.c.automine
        try:
            payload = json.dumps(a.input, separators=(",", ":")).encode()
            self.chain.submit_task(a.version, self.chain.address, a.model, int(a.fee), payload)
        except ChainError as e:
            log.error("Automine submitTask failed %s", e)
        if a.enabled:
            self.db.queue_job("automine", priority=5, waituntil=self.chain.now() + a.delay, concurrent=False, data={})

    def process_task(self, taskid: str, txid: str) -> None:
        """index.ts:506-564."""
        t = self.lookup_and_insert_task(taskid)
        if t is None:
            raise RuntimeError("could not look up task")
        if t.version != 0:
            log.debug("Task (%s) has invalid version %s", taskid, t.version)
            self.db.store_invalid_task(taskid)
            return
        fr = check_model_filter(self.models, model=t.model, now=self.chain.now(), fee=t.fee, blocktime=t.blocktime,
                                owner=t.owner)
        if not fr.model_enabled:
            log.debug("Task (%s) is using non-enabled Model (%s)", taskid, t.model)
            return
        if not fr.filter_passed:
            log.debug("Task (%s) does not pass filter", taskid)
            return
        if not self.lookup_and_insert_task_input(taskid, t.cid, txid, fr.model_template):
            return
        self.db.queue_job("solve", priority=20, waituntil=0, concurrent=False, data={"taskid": taskid})

    def get_cid(self, model: ModelEntry, taskid: str, inp: dict, gpus=None) -> Optional[str]:
        """``default__getcid`` (models.ts:34-54): files -> wrapped-directory CIDv0 -> 0x hex."""
        if self.c.evilmode:
            return EVIL_CID
        files = self._retry(lambda: self.ml.getfiles(model, taskid, inp, gpus=tuple(gpus) if gpus else None))
        if not files:
            log.error("task (%s) getfiles failed", taskid)
            return None
        cid58 = self._retry(lambda: ipfs.pin_files(self.c, taskid, files, pin=self.pin))
        if not cid58:
            log.error("task (%s) pinning failed", taskid)
            return None
        return "0x" + b58decode(cid58).hex()

    def process_solve(self, taskid: str, gpus=None) -> None:
        """index.ts:566-672."""
        t0 = time.monotonic()
        s = self._retry(lambda: self.chain.solutions(taskid))
        if s is not None and s.validator != ZERO_ADDR:
            return
        t = self._retry(lambda: self.lookup_and_insert_task(taskid))
        if not t:
            raise RuntimeError("could not look up task")
        m = get_model_by_id(self.models, t.model)
        if m is None:
            log.error("Task (%s) could not find model (%s)", taskid, t.model)
            return
        inp = self.db.get_task_input(taskid, t.cid)
        if inp is None:
            log.warning("Task (%s) input not found in db", taskid)
            return
        cid = self.get_cid(m, taskid, inp, gpus)
        if not cid:
            log.error("Task (%s) CID could not be generated", taskid)
            return
        t_model = time.monotonic()
        commitment = generate_commitment(self.chain.address, taskid, cid)

        def commit():
            try:
                self.chain.signal_commitment(commitment)
                return True
            except ChainError as e:
                if "nonce" in str(e).lower() or "underpriced" in str(e).lower():
                    raise                       # a sender race, not a protocol answer: retry, the GPU work is done
                log.error("Commitment submission failed %s", e)
                return False
        ok = self._retry(commit, 5, 1.25)
        if ok is None:
            raise RuntimeError(f"signalCommitment for {taskid} kept failing; the solve is kept for replay")
        if not ok:
            return

        def submit():
            try:
                self.chain.submit_solution(taskid, cid)
                self.db.queue_job("claim", priority=50, waituntil=self.chain.now() + CLAIM_DELAY, concurrent=False,
                                  data={"taskid": taskid})
                return True
            except ChainError as e:
                ex = self._retry(lambda: self.chain.solutions(taskid))
                if ex is None or ex.validator == ZERO_ADDR:
                    raise RuntimeError(f"unknown error submitting solution for {taskid} with cid {cid}: {e}")
                if ex.cid == cid:
                    log.info("Solution found for %s matches our cid %s", taskid, cid)
                    return True
                log.info("Solution found with cid %s does not match ours %s", ex.cid, cid)
                self.contest_solution(taskid)
                return True
        self._retry(submit, 3, 1.25)
        self.solved.append({"taskid": taskid, "cid": cid, "model_ms": (t_model - t0) * 1e3,
                            "total_ms": (time.monotonic() - t0) * 1e3, "gpus": gpus})

    def contest_solution(self, taskid: str) -> None:
        """index.ts:674-707."""
        try:
            self.chain.submit_contestation(taskid)
            self.db.queue_job("contestationVoteFinish", priority=30, waituntil=self.chain.now() + CONTESTATION_FINISH_DELAY,
                              concurrent=False, data={"taskid": taskid})
        except ChainError as e:
            log.warning("Failed to contest %s (%s): voting instead", taskid, e)
            self.vote_on_contestation(taskid, True)

    def vote_on_contestation(self, taskid: str, yea: bool) -> None:
        """index.ts:709-726."""
        if self._retry(lambda: self.chain.contestation_voted(taskid, self.chain.address)):
            log.debug("Contestation %s already voted", taskid)
            return
        try:
            self.chain.vote_on_contestation(taskid, yea)
        except ChainError as e:
            log.error("Failed voting on contestation %s: %s", taskid, e)

    def process_contestation_vote_finish(self, taskid: Optional[str]) -> None:
        """The reference leaves this a stub (index.ts:392-395); here stakes really get released."""
        if not taskid:
            return
        c = self.chain.contestations(taskid)
        if c.validator == ZERO_ADDR:
            return
        try:
            self.chain.contestation_vote_finish(taskid, 64)
        except ChainError as e:
            log.warning("contestationVoteFinish %s failed: %s", taskid, e)

    def process_claim(self, taskid: str) -> None:
        """index.ts:728-750."""
        s = self._retry(lambda: self.chain.solutions(taskid))
        if s is None or s.claimed:
            log.warning("Solution (%s) already claimed", taskid)
            return
        self.chain.claim_solution(taskid)
        log.info("Claim %s", taskid)

    # ------------------------------------------------------------------ dispatcher (index.ts:879-958)
    def _run_job(self, job: Job, gpus=None) -> None:
        d = job.decoded()
        m = job.method
        if m == "automine":
            self.process_automine()
        elif m == "validatorStake":
            self.process_validator_stake()
        elif m == "task":
            self.process_task(d["taskid"], d.get("txid", ""))
        elif m == "solve":
            self.process_solve(d["taskid"], gpus)
        elif m == "claim":
            self.process_claim(d["taskid"])
        elif m == "pinTaskInput":
            self.process_pin_task_input(d["taskid"], d["input"])
        elif m == "contestationVoteFinish":
            self.process_contestation_vote_finish(d.get("taskid"))
        else:
            log.error("Job (%s) method (%s) has no implementation", job.id, m)
            raise FatalMinerError(1)

    def _finish(self, job: Job, err: Optional[BaseException]) -> None:
        if err is not None:
            if isinstance(err, FatalMinerError):
                raise err
            log.error("Job (%s) [%s] failed: %s", job.id, job.method, err)
            d = job.decoded()
            tries = int(d.get("_retries", 0))
            if job.method == "solve" and tries < self.solve_retries:
                # the reference dead-letters and forgets (index.ts:935-937); a solve is worth a bounded replay
                d["_retries"] = tries + 1
                self.db.queue_job("solve", priority=job.priority, waituntil=self.chain.now() + 5 * (tries + 1),
                                  concurrent=False, data=d)
            else:
                self.db.store_failed_job(job)
        self.db.delete_job(job.id)
        with self._lock:
            self._inflight.pop(job.id, None)

    def process_jobs(self, jobs: Sequence[Job]) -> int:
        """Runs every due job: solves go to the GPU pool (asynchronously), everything else runs inline."""
        started = 0
        tnow = self.chain.now()
        for job in jobs:
            with self._lock:
                if job.id in self._inflight or job.waituntil > tnow:
                    continue
            started += 1
            if job.method == "solve":
                d = job.decoded()
                task = self.db.get_task(d["taskid"]) or {}
                mid = str(task.get("modelid", "unknown")).lower()
                # the cost model is keyed by model family (template name); unknown ids keep their id as the key
                family = next((m.name for m in self.models if m.id.lower() == mid and m.name), mid)
                fut = self.pool.submit(family, lambda gpus, j=job: self._run_job(j, gpus), priority=job.priority,
                                       age_key=float(task.get("blocktime") or 0))
                with self._lock:
                    self._inflight[job.id] = fut
                fut.add_done_callback(lambda f, j=job: self._finish(j, f.exception()))
            else:
                err = None
                try:
                    self._run_job(job)
                except FatalMinerError:
                    raise
                except BaseException as e:  # noqa: BLE001
                    err = e
                self._finish(job, err)
        return started

    # ------------------------------------------------------------------ boot (index.ts:960-1101)
    def version_check(self) -> None:
        v = self._retry(self.chain.version)
        if v is None or v > 0:
            log.error("version mismatch, have miner version 0 and chain is %s. Upgrade your miner", v)
            raise FatalMinerError(1)

    def self_test(self, model: ModelEntry, inp: dict, expected_cid: Optional[str]) -> str:
        """Boot self-test (index.ts:983-1001): fixed prompt + seed must hash to the golden CID."""
        cid = self.get_cid(model, "startup-test-taskid", inp)
        if expected_cid is not None and cid != expected_cid:
            log.error("Model reproducibility test failed: %s != %s", cid, expected_cid)
            raise FatalMinerError(1)
        return cid

    def boot_self_test(self, write: bool = False) -> Dict[str, str]:
        """index.ts:983-1001 for every enabled model: the fixed probe must hash to the golden recorded for this
        (template, library build, GPU model). A mismatch is fatal; a model without a golden is recorded when
        ``write`` (the ``golden --write`` CLI task) and reported otherwise."""
        from .golden import environment_key, selftest_input
        out: Dict[str, str] = {}
        if self.goldens is None or self.c.evilmode:
            return out
        for m in self.models:
            if not m.mineable:
                continue
            key = environment_key(m.name or m.id)
            expected = None if write else self.goldens.get(key)
            if expected is None and not write:
                log.warning("self-test: no golden CID recorded for %s — run `golden --write` on a verified box", key)
                continue
            cid = self.self_test(m, selftest_input(m.template), expected)
            if cid is None:
                log.error("self-test: model %s produced no CID", key)
                raise FatalMinerError(1)
            if write:
                self.goldens.put(key, cid)
            out[key] = cid
        return out

    def bootup(self) -> None:
        self.db.clear_jobs_by_method("validatorStake")
        self.db.clear_jobs_by_method("automine")
        self.version_check()
        self.boot_self_test()
        self.db.queue_job("validatorStake", priority=30, waituntil=0, concurrent=False, data={})
        if self.c.automine.enabled:
            self.db.queue_job("automine", priority=5, waituntil=0, concurrent=False, data={})
        self.chain.subscribe(self.on_event)

    def step(self) -> int:
        self.chain.poll_events()
        return self.process_jobs(self.db.get_jobs())

    def run_forever(self) -> None:
        self.bootup()
        while not self.stopped:
            if self.step() == 0:
                self._sleep(self.c.scheduler.poll_ms / 1000.0)

    def run_until_idle(self, max_iters: int = 1000) -> None:
        """Test/driver helper: process until no job is due and nothing is in flight."""
        for _ in range(max_iters):
            n = self.step()
            self.pool.drain(timeout=60)
            with self._lock:
                busy = bool(self._inflight)
            if n == 0 and not busy:
                due = [j for j in self.db.get_jobs() if j.waituntil <= self.chain.now()]
                if not due:
                    return
