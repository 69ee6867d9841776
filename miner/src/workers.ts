// Worker registry and concurrent solve pump for the Arbius miner (the executing half of scheduler.ts).
//
// MiningConfig stays loadable unchanged: `ml.cog[modelid]` may be the stock `{ url }` or, additionally,
// `{ urls: [...], gpus: [[0], [1], ..., [0, 1], ...], family, costs_ms: { "1": 127, "2": 110 }, slots }` — one Cog
// endpoint per GPU subset (`python -m arbius_b200.worker.server --device cuda:N --port ...`). With `{ url }` alone the
// registry has one single-GPU worker and the miner behaves exactly as before.
//
// What changes in index.ts is small (see ../PATCH_index_ts.md): the Kandinsky2 `getfiles` calls `solveOnWorker`
// instead of posting to the one URL, and `processJobs` hands `solve` jobs to `SolvePump` instead of awaiting them one by
// one. Every task writes to its own directory cache/<taskid>/, because with several solves in flight the stock
// cache/out-1.png would be overwritten under the IPFS add.
import axios from 'axios';
import * as fs from 'fs';
import * as path from 'path';
import { CostModel, CostTable, Policy, SubsetPlanner } from './scheduler';

export interface CogEntry {
  url?: string;
  urls?: string[];
  gpus?: number[][];                 // GPU ids behind urls[i]; default: worker i drives GPU i alone
  family?: string;                   // cost-model key; default: the model id
  costs_ms?: { [p: string]: number };
  slots?: number;                    // tasks a single-GPU worker may batch
}

export interface Worker { url: string; gpus: number[]; }

function sameSet(a: number[], b: number[]): boolean {
  return a.length === b.length && a.every((g) => b.includes(g));
}

export class WorkerRegistry {
  workers: Worker[];
  family: string;
  planner: SubsetPlanner;
  cost: CostModel;

  constructor(modelid: string, entry: CogEntry) {
    const urls = entry.urls && entry.urls.length > 0 ? entry.urls : [entry.url as string];
    if (!urls[0]) throw new Error(`ml.cog[${modelid}] needs url or urls`);
    this.workers = urls.map((url, i) => ({ url, gpus: entry.gpus && entry.gpus[i] ? entry.gpus[i] : [i] }));
    this.family = entry.family || modelid;
    const singles = this.workers.filter((w) => w.gpus.length === 1).length;
    const n = [8, 4, 2, 1].find((p) => p <= Math.max(1, singles)) as number;
    this.planner = new SubsetPlanner(n, entry.slots || 1);
    const table: CostTable = {};
    if (entry.costs_ms) {
      table[this.family] = {};
      for (const k of Object.keys(entry.costs_ms)) table[this.family][Number(k)] = entry.costs_ms[k];
    }
    this.cost = new CostModel(table);
  }

  // the endpoint that drives exactly this GPU subset (a wider task needs a worker that was started on that subset)
  pickWorker(gpus: number[]): Worker | null {
    return this.workers.find((w) => sameSet(w.gpus, gpus)) || null;
  }

  // subset sizes some worker can actually serve: the planner must not hand out a subset nobody listens on
  servedSizes(): number[] {
    return [1, 2, 4, 8].filter((p) => this.workers.some((w) => w.gpus.length === p));
  }
}

export function taskCacheDir(cacheRoot: string, taskid: string): string {
  const dir = path.join(cacheRoot, taskid);
  fs.mkdirSync(dir, { recursive: true });
  return dir;
}

// POST {input} to the worker and write the single data-URI output to cache/<taskid>/<filename> (the Cog contract the
// stock getfiles speaks: one element, `data:<type>/<subtype>;base64,` prefix)
export async function solveOnWorker(worker: Worker, cacheRoot: string, taskid: string, input: any, filename: string): Promise<string[]> {
  const res = await axios.post(worker.url, { input });
  if (!res || !res.data || !res.data.output || res.data.output.length != 1) {
    throw new Error('unable to getfiles -- data.output length not 1');
  }
  const data = String(res.data.output[0]).replace(/^data:\w+\/\w+;base64,/, '');
  const out = path.join(taskCacheDir(cacheRoot, taskid), filename);
  fs.writeFileSync(out, Buffer.from(data, 'base64'));
  return [path.join(taskid, filename)];
}

interface QueuedSolve { taskid: string; priority: number; age: number; seq: number; run: (w: Worker) => Promise<void>; done: () => void; fail: (e: any) => void; }

// Keeps up to one solve per free GPU subset in flight. `submit` resolves when the task's solve has finished (or
// rejects with its error), so processJobs can delete / dead-letter the job row afterwards — a crash between the two no
// longer loses the job, which the stock "delete before completion" order does.
export class SolvePump {
  private queue: QueuedSolve[] = [];
  private seq = 0;
  inflight = 0;
  completed: Array<{ taskid: string; p: number; ms: number }> = [];

  constructor(private reg: WorkerRegistry, private policy: Policy = 'throughput') {}

  submit(taskid: string, priority: number, age: number, run: (w: Worker) => Promise<void>): Promise<void> {
    return new Promise<void>((done, fail) => {
      this.queue.push({ taskid, priority, age, seq: this.seq++, run, done, fail });
      this.queue.sort((a, b) => (b.priority - a.priority) || (a.age - b.age) || (a.seq - b.seq));
      this.dispatch();
    });
  }

  private dispatch(): void {
    while (this.queue.length > 0) {
      const item = this.queue[0];
      let p = this.reg.planner.chooseP(this.reg.family, this.queue.length, this.reg.cost, this.policy);
      if (p === 0) return;
      const served = this.reg.servedSizes();
      while (p > 1 && !served.includes(p)) p = p / 2;
      const gpus = this.reg.planner.acquire(p);
      if (!gpus) return;
      const worker = this.reg.pickWorker(gpus);
      if (!worker) { this.reg.planner.release(gpus); return; }
      this.queue.shift();
      this.inflight += 1;
      const t0 = Date.now();
      item.run(worker).then(item.done, item.fail).finally(() => {
        const ms = Date.now() - t0;
        this.reg.planner.release(gpus);
        this.inflight -= 1;
        this.reg.cost.update(this.reg.family, gpus.length, ms);
        this.completed.push({ taskid: item.taskid, p: gpus.length, ms });
        this.dispatch();
      });
    }
  }
}
