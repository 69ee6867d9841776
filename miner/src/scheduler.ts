// Multi-GPU task scheduler for the Arbius miner: packs independent on-chain tasks onto GPU subsets.
//
// The stock miner awaits one `solve` job at a time (index.ts: `concurrent: false`, the second loop of processJobs),
// posts it to the single Cog URL of its model and writes the answer to the fixed path cache/out-1.png. On a box with
// eight GPUs that leaves seven idle. This module is the planning half of the replacement: pure functions and small
// classes with no I/O, so the same decisions can be replayed in tests (test/scheduler.test.ts) and are mirrored one to
// one by arbius_b200/miner/scheduler.py (the in-process Python miner); both are pinned to the same golden traces
// (tests/data/scheduler_traces.json).
//
//   * CostModel      measured milliseconds per (model family, GPU count P); the split kernels of the B200 workers
//                    are bit-identical for P in {1,2,4,8}, so P is purely a performance knob
//   * SubsetPlanner  which GPUs are free; hands out aligned power-of-two subsets (NVSwitch peers of a subset are
//                    always the same ranks, so a worker per subset can map its peers once)
//   * policy         "throughput": the P that maximises tasks/hour/GPU while work is queued;
//                    "latency": fewer tasks than free GPUs -> widen a task to the largest subset that still
//                    shortens it (first valid solution wins: EngineV1.submitSolution)
//   * older tasks first inside a priority class (paper: older tasks are more winnable)

export const SUBSET_SIZES = [1, 2, 4, 8] as const;
export type Policy = 'throughput' | 'latency';
export type CostTable = { [family: string]: { [p: number]: number } };

function bitLength(n: number): number {
  let b = 0;
  while (n > 0) { b += 1; n = Math.floor(n / 2); }
  return b;
}

export class CostModel {
  table: CostTable;
  efficiency: number; // assumed parallel efficiency per doubling when unmeasured: 0.5 = "no speed-up"

  constructor(table: CostTable = {}, efficiency = 0.5) {
    this.table = table;
    this.efficiency = efficiency;
  }

  ms(family: string, p: number): number {
    const t = this.table[family];
    if (!t || Object.keys(t).length === 0) {
      return 1000.0 / Math.pow(2 * this.efficiency, bitLength(p) - 1);
    }
    if (t[p] !== undefined) return t[p];
    const keys = Object.keys(t).map(Number);
    const below = keys.filter((k) => k <= p);
    const known = below.length > 0 ? Math.max(...below) : Math.min(...keys);
    const doublings = bitLength(Math.floor(p / known)) - 1;
    return t[known] / Math.pow(2 * this.efficiency, doublings);
  }

  update(family: string, p: number, ms: number, alpha = 0.2): void {
    if (!this.table[family]) this.table[family] = {};
    const cur = this.table[family][p];
    this.table[family][p] = cur === undefined ? ms : (1 - alpha) * cur + alpha * ms;
  }

  // P maximising tasks/hour per GPU = 1 / (ms(P) * P); ties go to the smaller P
  bestThroughputP(family: string, maxP: number): number {
    let best = 1;
    let bestCost = Infinity;
    for (const p of SUBSET_SIZES) {
      if (p > maxP) continue;
      const c = this.ms(family, p) * p;
      if (c < bestCost) { bestCost = c; best = p; }
    }
    return best;
  }

  // largest P whose latency is at least minGain x better than the best smaller one
  bestLatencyP(family: string, maxP: number, minGain = 1.15): number {
    let best = 1;
    for (const p of SUBSET_SIZES) {
      if (p === 1 || p > maxP) continue;
      if (this.ms(family, best) / this.ms(family, p) >= minGain) best = p;
    }
    return best;
  }
}

export class SubsetPlanner {
  n: number;
  slots: number; // single-GPU tasks that may share a GPU (the worker batches them into one graph replay)
  load: number[];

  constructor(nGpus: number, slots = 1) {
    if (!(SUBSET_SIZES as readonly number[]).includes(nGpus)) throw new Error('nGpus must be 1, 2, 4 or 8');
    this.n = nGpus;
    this.slots = Math.max(1, Math.floor(slots));
    this.load = new Array(nGpus).fill(0);
  }

  freeCount(): number {
    return this.load.reduce((acc, c) => acc + (this.slots - c), 0);
  }

  private idle(start: number, p: number): boolean {
    for (let i = start; i < start + p; i++) if (this.load[i] !== 0) return false;
    return true;
  }

  acquire(p: number): number[] | null {
    if (p === 1) {
      let g = 0; // spread first, then fill second slots
      for (let i = 1; i < this.n; i++) if (this.load[i] < this.load[g]) g = i;
      if (this.load[g] >= this.slots) return null;
      this.load[g] += 1;
      return [g];
    }
    for (let start = 0; start < this.n; start += p) {
      if (this.idle(start, p)) {
        const out: number[] = [];
        for (let i = start; i < start + p; i++) { this.load[i] = this.slots; out.push(i); }
        return out;
      }
    }
    return null;
  }

  largestFreeAligned(): number {
    for (const p of [...SUBSET_SIZES].reverse()) {
      if (p > this.n) continue;
      if (p === 1) return Math.min(...this.load) < this.slots ? 1 : 0;
      for (let start = 0; start < this.n; start += p) if (this.idle(start, p)) return p;
    }
    return 0;
  }

  release(gpus: number[]): void {
    if (gpus.length === 1) { this.load[gpus[0]] -= 1; return; }
    for (const i of gpus) this.load[i] = 0;
  }

  // how many GPUs the next task gets, given `waiting` queued tasks (including it); 0 = nothing free
  chooseP(family: string, waiting: number, cost: CostModel, policy: Policy): number {
    const avail = this.largestFreeAligned();
    if (avail === 0) return 0;
    if (policy === 'throughput' && waiting >= this.freeCount()) {
      return Math.min(cost.bestThroughputP(family, avail), avail);
    }
    const idleGpus = this.load.filter((c) => c === 0).length;
    const share = Math.max(1, Math.floor(idleGpus / Math.max(waiting, 1)));
    let cap = 1;
    for (const p of SUBSET_SIZES) if (p <= share && p <= avail) cap = p;
    return cost.bestLatencyP(family, cap);
  }
}

export interface SimResult { tasksPerHour: number; p50LatencyS: number; makespanS: number; assignments: number[][]; }

// Event-driven replay of (arrival seconds, family) traces against the planner; `assignments` lists the GPUs every
// task got, in dispatch order (the golden traces pin these).
export function simulate(tasks: Array<[number, string]>, nGpus: number, cost: CostModel, policy: Policy): SimResult {
  const planner = new SubsetPlanner(nGpus);
  const arrivals = [...tasks].sort((a, b) => (a[0] - b[0]) || (a[1] < b[1] ? -1 : a[1] > b[1] ? 1 : 0));
  const pending: Array<[number, string]> = [];
  let running: Array<[number, number[]]> = [];
  const lat: number[] = [];
  const assignments: number[][] = [];
  let now = 0.0;
  let i = 0;
  while (i < arrivals.length || pending.length > 0 || running.length > 0) {
    while (i < arrivals.length && arrivals[i][0] <= now) { pending.push(arrivals[i]); i += 1; }
    let progressed = true;
    while (pending.length > 0 && progressed) {
      progressed = false;
      const [arr, fam] = pending[0];
      const p = planner.chooseP(fam, pending.length, cost, policy);
      if (p > 0) {
        const gpus = planner.acquire(p);
        if (gpus) {
          pending.shift();
          const end = now + cost.ms(fam, p) / 1e3;
          running.push([end, gpus]);
          running.sort((a, b) => a[0] - b[0]);
          lat.push(end - arr);
          assignments.push(gpus);
          progressed = true;
        }
      }
    }
    const nxt: number[] = [];
    if (running.length > 0) nxt.push(running[0][0]);
    if (i < arrivals.length) nxt.push(arrivals[i][0]);
    if (nxt.length === 0) break;
    now = Math.max(now, Math.min(...nxt));
    while (running.length > 0 && running[0][0] <= now) {
      const done = running.shift()!;
      planner.release(done[1]);
    }
  }
  lat.sort((a, b) => a - b);
  return {
    tasksPerHour: lat.length / Math.max(now, 1e-9) * 3600.0,
    p50LatencyS: lat.length > 0 ? lat[Math.floor(lat.length / 2)] : 0.0,
    makespanS: now,
    assignments,
  };
}
