"""Multi-GPU task scheduler: packs independent on-chain tasks onto GPU subsets.

The reference awaits one ``solve`` at a time (reference: miner/src/index.ts:555-563 ``concurrent:false``,
:920-957) — one GPU request in flight, fixed output path — which caps a box at one task at a time.
This scheduler keeps every GPU busy and decides, per task, how many GPUs to give it:

  * cost model: measured milliseconds per (model family, GPU count P) — ``CostModel``; the split kernels are
    bit-identical for P in {1,2,4,8}, so P is purely a performance knob;
  * policy "throughput": P that maximises tasks/hour/GPU (almost always 1) while work is queued;
  * policy "latency": when fewer tasks than GPUs are waiting, widen tasks to the largest aligned subset that
    still shortens their latency (first valid solution wins: reference paper/paper.tex:41,
    contract/contracts/EngineV1.sol:790-794);
  * older tasks first inside a priority class (paper/paper.tex:79-81: older tasks are more winnable).

``SubsetPlanner`` is pure logic (simulated in tests); ``GpuPool`` executes plans with worker threads and
per-subset predictors. Subsets are aligned, power-of-two slices of the GPU list, so NVSwitch peers of a
subset are always the same ranks and peer-memory mappings can be set up once.
"""
from __future__ import annotations

import heapq
import threading
import time
from concurrent.futures import Future
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

SUBSET_SIZES = (1, 2, 4, 8)


@dataclass
class CostModel:
    """ms per task by family and GPU count. Missing P entries are extrapolated with ``efficiency``."""
    table: Dict[str, Dict[int, float]] = field(default_factory=dict)
    efficiency: float = 0.5            # assumed parallel efficiency per doubling when unmeasured:
    #                                    0.5 = "no speed-up" — tasks are only widened on measured evidence

    # ms per task measured on B200 at the template defaults (profiles/multigpu_n8_v1.md); the P=2 rows are
    # the guidance-parallel pair predictors and only count where such a predictor was actually built
    # ms per task, round-2 end state (profiles/bench_r2_n1_v2.json, bench_r2_n8_v0.json latency_mode); the two-GPU
    # guidance-parallel entries of ZeroScope / Kandinsky are the round-1 ratios applied to the new one-GPU times
    MEASURED_B200 = {"anythingv3": {1: 116.0, 2: 98.0}, "zeroscopev2xl": {1: 13100.0, 2: 6800.0},
                     "kandinsky2": {1: 780.0, 2: 620.0}, "damo": {1: 1135.0}}

    # ms per task of the frame-parallel text-to-video predictors (tests/mp_frame_parallel.py scaled to 50 steps)
    MEASURED_B200_FRAMES = {"zeroscopev2xl": {2: 7900.0, 4: 4566.0, 8: 2689.0}, "damo": {2: 993.0}}   # profiles/multigpu_frame_parallel_v0.md, bench_r2_n8_v0.json

    @classmethod
    def measured_b200(cls, pair_families: Sequence[str] = (), frame_families: Sequence[str] = (),
                      frame_sizes: Sequence[int] = ()) -> "CostModel":
        table = {}
        for fam, row in cls.MEASURED_B200.items():
            table[fam] = {p: ms for p, ms in row.items() if p == 1 or fam in pair_families}
        for fam in frame_families:
            for p, ms in cls.MEASURED_B200_FRAMES.get(fam, {}).items():
                if p in frame_sizes:
                    table.setdefault(fam, {})[p] = ms
        return cls(table=table)

    def ms(self, family: str, p: int) -> float:
        t = self.table.get(family)
        if not t:
            return 1000.0 / ((2 * self.efficiency) ** (p.bit_length() - 1))
        if p in t:
            return t[p]
        known = max(k for k in t if k <= p) if any(k <= p for k in t) else min(t)
        doublings = (p // known).bit_length() - 1
        return t[known] / ((2 * self.efficiency) ** doublings)

    def update(self, family: str, p: int, ms: float, alpha: float = 0.2) -> None:
        cur = self.table.setdefault(family, {}).get(p)
        self.table[family][p] = ms if cur is None else (1 - alpha) * cur + alpha * ms

    def best_throughput_p(self, family: str, max_p: int) -> int:
        """P maximising tasks/hour per GPU = 1 / (ms(P) * P)."""
        return min((p for p in SUBSET_SIZES if p <= max_p), key=lambda p: self.ms(family, p) * p)

    def best_latency_p(self, family: str, max_p: int, min_gain: float = 1.15) -> int:
        """Largest P whose latency is at least ``min_gain`` x better than the next smaller one."""
        best = 1
        for p in SUBSET_SIZES:
            if p == 1 or p > max_p:
                continue
            if self.ms(family, best) / self.ms(family, p) >= min_gain:
                best = p
        return best


@dataclass(order=True)
class _Queued:
    sort_key: Tuple[int, float, int]
    family: str = field(compare=False)
    fn: Callable = field(compare=False)
    future: Future = field(compare=False)
    submitted: float = field(compare=False, default=0.0)


class SubsetPlanner:
    """Tracks which GPUs are free and hands out aligned power-of-two subsets. ``slots`` > 1 lets that many single-GPU
    tasks share a GPU (the predictor batches them into one graph replay); wider subsets need wholly idle GPUs."""

    def __init__(self, n_gpus: int, slots: int = 1):
        if n_gpus not in SUBSET_SIZES:
            raise ValueError("n_gpus must be 1, 2, 4 or 8")
        self.n = n_gpus
        self.slots = max(1, int(slots))
        self.load = [0] * n_gpus

    @property
    def busy(self) -> List[bool]:
        return [c > 0 for c in self.load]

    def free_count(self) -> int:
        """Free task slots (= idle GPUs when slots == 1)."""
        return sum(self.slots - c for c in self.load)

    def acquire(self, p: int) -> Optional[Tuple[int, ...]]:
        if p == 1:
            g = min(range(self.n), key=lambda i: (self.load[i], i))      # spread first, then fill second slots
            if self.load[g] >= self.slots:
                return None
            self.load[g] += 1
            return (g,)
        for start in range(0, self.n, p):
            if not any(self.load[start:start + p]):
                for i in range(start, start + p):
                    self.load[i] = self.slots
                return tuple(range(start, start + p))
        return None

    def largest_free_aligned(self) -> int:
        for p in reversed(SUBSET_SIZES):
            if p > self.n:
                continue
            if p == 1:
                return 1 if min(self.load) < self.slots else 0
            for start in range(0, self.n, p):
                if not any(self.load[start:start + p]):
                    return p
        return 0

    def release(self, gpus: Sequence[int]) -> None:
        if len(gpus) == 1:
            self.load[gpus[0]] -= 1
            return
        for i in gpus:
            self.load[i] = 0

    def choose_p(self, family: str, waiting: int, cost: CostModel, policy: str) -> int:
        """How many GPUs the next task gets, given ``waiting`` queued tasks (including it)."""
        avail = self.largest_free_aligned()
        if avail == 0:
            return 0
        if policy == "throughput" and waiting >= self.free_count():
            return min(cost.best_throughput_p(family, avail), avail)
        # fewer tasks than free GPUs (or latency policy): spread free GPUs over the waiting tasks
        share = max(1, self.load.count(0) // max(waiting, 1))
        cap = 1
        for p in SUBSET_SIZES:
            if p <= share and p <= avail:
                cap = p
        return cost.best_latency_p(family, cap)


class GpuPool:
    """Runs queued tasks on GPU subsets. ``runner(family, gpus, fn)`` executes ``fn(gpus)`` wherever the subset's
    predictor lives (thread here; a worker process per GPU in production)."""

    def __init__(self, n_gpus: int, cost: Optional[CostModel] = None, policy: str = "throughput",
                 clock: Callable[[], float] = time.monotonic, slots: int = 1):
        self.planner = SubsetPlanner(n_gpus, slots)
        self.cost = cost or CostModel()
        self.policy = policy
        self._clock = clock
        self._q: List[_Queued] = []
        self._seq = 0
        self._cv = threading.Condition()
        self._inflight = 0
        self.completed: List[Tuple[str, int, float]] = []      # (family, P, ms)

    def submit(self, family: str, fn: Callable[[Tuple[int, ...]], object], priority: int = 0,
               age_key: float = 0.0) -> Future:
        fut: Future = Future()
        with self._cv:
            self._seq += 1
            heapq.heappush(self._q, _Queued((-priority, age_key, self._seq), family, fn, fut, self._clock()))
            self._dispatch_locked()
        return fut

    def _dispatch_locked(self) -> None:
        while self._q:
            item = self._q[0]
            p = self.planner.choose_p(item.family, len(self._q), self.cost, self.policy)
            if p == 0:
                return
            gpus = self.planner.acquire(p)
            if gpus is None:
                return
            heapq.heappop(self._q)
            self._inflight += 1
            threading.Thread(target=self._run, args=(item, gpus), daemon=True).start()

    def _run(self, item: _Queued, gpus: Tuple[int, ...]) -> None:
        t0 = self._clock()
        try:
            res = item.fn(gpus)
            item.future.set_result(res)
        except BaseException as e:  # noqa: BLE001
            item.future.set_exception(e)
        finally:
            ms = (self._clock() - t0) * 1e3
            with self._cv:
                self.planner.release(gpus)
                self._inflight -= 1
                self.cost.update(item.family, len(gpus), ms)
                self.completed.append((item.family, len(gpus), ms))
                self._dispatch_locked()
                self._cv.notify_all()

    def drain(self, timeout: Optional[float] = None) -> bool:
        with self._cv:
            return self._cv.wait_for(lambda: not self._q and self._inflight == 0, timeout)

    @property
    def inflight(self) -> int:
        return self._inflight


def simulate(tasks: Sequence[Tuple[float, str]], n_gpus: int, cost: CostModel, policy: str) -> Dict[str, float]:
    """Event-driven replay of (arrival_s, family) traces against the planner: tasks/hour and p50 latency."""
    planner = SubsetPlanner(n_gpus)
    pending: List[Tuple[float, int, str]] = []
    running: List[Tuple[float, Tuple[int, ...]]] = []
    lat: List[float] = []
    assignments: List[List[int]] = []
    now, i, seq = 0.0, 0, 0
    arrivals = sorted(tasks)
    while i < len(arrivals) or pending or running:
        # admit arrivals up to now
        while i < len(arrivals) and arrivals[i][0] <= now:
            pending.append((arrivals[i][0], seq, arrivals[i][1]))
            seq += 1
            i += 1
        # dispatch
        progressed = True
        while pending and progressed:
            progressed = False
            arr, _, fam = pending[0]
            p = planner.choose_p(fam, len(pending), cost, policy)
            if p:
                gpus = planner.acquire(p)
                if gpus:
                    pending.pop(0)
                    end = now + cost.ms(fam, p) / 1e3
                    heapq.heappush(running, (end, gpus))
                    lat.append(end - arr)
                    assignments.append(list(gpus))
                    progressed = True
        # advance time
        nxt = []
        if running:
            nxt.append(running[0][0])
        if i < len(arrivals):
            nxt.append(arrivals[i][0])
        if not nxt:
            break
        now = max(now, min(nxt))
        while running and running[0][0] <= now:
            _, gpus = heapq.heappop(running)
            planner.release(gpus)
    lat.sort()
    return {"tasks_per_hour": len(lat) / max(now, 1e-9) * 3600.0, "p50_latency_s": lat[len(lat) // 2] if lat else 0.0,
            "makespan_s": now, "assignments": assignments}
