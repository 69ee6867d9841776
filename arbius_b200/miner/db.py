"""Durable miner state: SQLite tables and the priority job queue.

Same tables, columns and semantics as the reference (reference: miner/src/sql/*.sql, miner/src/db.ts:24-378)
so an existing ``db.sqlite`` can be opened unchanged. Differences, all deliberate:
  * ``invalid_tasks`` IS created (the reference forgets to load invalid_tasks.sql — SURVEY Appendix B.4);
  * a ``running`` lease column is *not* added to ``jobs``: in-flight jobs are tracked in memory and the row
    is deleted only after the job finished, so a crash never loses a job;
  * one connection guarded by a lock, usable from the scheduler's worker threads.
"""
from __future__ import annotations

import json
import sqlite3
import threading
from dataclasses import dataclass
from typing import Any, Dict, List, Optional

from ..protocol.seed import taskid2seed

SCHEMA = """
CREATE TABLE IF NOT EXISTS tasks (id TEXT PRIMARY KEY, modelid TEXT, fee TEXT, address TEXT, blocktime TEXT,
    version INT, cid TEXT, retracted BOOLEAN DEFAULT FALSE);
CREATE INDEX IF NOT EXISTS tasks_id ON tasks(id);
CREATE INDEX IF NOT EXISTS tasks_modelid ON tasks(modelid);
CREATE INDEX IF NOT EXISTS tasks_address ON tasks(address);
CREATE TABLE IF NOT EXISTS task_inputs (taskid TEXT, cid TEXT, data TEXT);
CREATE INDEX IF NOT EXISTS task_inputs_taskid ON task_inputs(taskid);
CREATE INDEX IF NOT EXISTS task_inputs_cid ON task_inputs(cid);
CREATE TABLE IF NOT EXISTS solutions (id INTEGER PRIMARY KEY AUTOINCREMENT, taskid TEXT, validator TEXT,
    blocktime TEXT, claimed BOOLEAN, cid TEXT);
CREATE INDEX IF NOT EXISTS solutions_taskid ON solutions(taskid);
CREATE INDEX IF NOT EXISTS solutions_validator ON solutions(validator);
CREATE TABLE IF NOT EXISTS contestations (id INTEGER PRIMARY KEY AUTOINCREMENT, taskid TEXT, validator TEXT,
    blocktime TEXT, finish_start_index TEXT);
CREATE INDEX IF NOT EXISTS contestations_taskid ON contestations(taskid);
CREATE INDEX IF NOT EXISTS contestations_validator ON contestations(validator);
CREATE TABLE IF NOT EXISTS contestation_votes (id INTEGER PRIMARY KEY AUTOINCREMENT, taskid TEXT, validator TEXT,
    yea BOOLEAN);
CREATE INDEX IF NOT EXISTS contestation_votes_taskid ON contestation_votes(taskid);
CREATE INDEX IF NOT EXISTS contestation_votes_validator ON contestation_votes(validator);
CREATE TABLE IF NOT EXISTS jobs (id INTEGER PRIMARY KEY AUTOINCREMENT, priority INTEGER, waituntil INTEGER,
    concurrent BOOLEAN, method TEXT, data TEXT);
CREATE TABLE IF NOT EXISTS failed_jobs (id INTEGER PRIMARY KEY AUTOINCREMENT, method TEXT, data TEXT);
CREATE TABLE IF NOT EXISTS models (id TEXT PRIMARY KEY, addr TEXT, mineable BOOLEAN, cid TEXT);
CREATE TABLE IF NOT EXISTS invalid_tasks (taskid TEXT PRIMARY KEY);
CREATE INDEX IF NOT EXISTS invalid_tasks_taskid ON invalid_tasks(taskid);
"""

JOB_METHODS = ("automine", "validatorStake", "task", "solve", "claim", "pinTaskInput", "contestationVoteFinish")


@dataclass
class Job:
    id: int
    priority: int
    waituntil: int
    concurrent: bool
    method: str
    data: str

    def decoded(self) -> dict:
        return json.loads(self.data) if self.data else {}


class MinerDB:
    def __init__(self, path: str = ":memory:"):
        self._lock = threading.RLock()
        self._db = sqlite3.connect(path, check_same_thread=False, isolation_level=None)
        self._db.row_factory = sqlite3.Row
        with self._lock:
            self._db.executescript(SCHEMA)

    def close(self) -> None:
        with self._lock:
            self._db.close()

    def _exec(self, sql: str, args=()) -> sqlite3.Cursor:
        with self._lock:
            return self._db.execute(sql, args)

    def _one(self, sql: str, args=()) -> Optional[sqlite3.Row]:
        with self._lock:
            return self._db.execute(sql, args).fetchone()

    def _all(self, sql: str, args=()) -> List[sqlite3.Row]:
        with self._lock:
            return self._db.execute(sql, args).fetchall()

    # ---------------------------------------------------------------- tasks
    def store_task(self, taskid: str, modelid: str, fee: int, owner: str, blocktime: int, version: int, cid: str):
        """db.ts dbStoreTask: INSERT OR IGNORE."""
        self._exec("INSERT OR IGNORE INTO tasks (id, modelid, fee, address, blocktime, version, cid) VALUES (?,?,?,?,?,?,?)",
                   (taskid, modelid, str(fee), owner, str(blocktime), version, cid))

    def get_task(self, taskid: str) -> Optional[dict]:
        r = self._one("SELECT * FROM tasks WHERE id=?", (taskid,))
        return dict(r) if r else None

    def update_task_set_retracted(self, taskid: str) -> None:
        self._exec("UPDATE tasks SET retracted=true WHERE id=?", (taskid,))

    def store_invalid_task(self, taskid: str) -> None:
        self._exec("INSERT OR IGNORE INTO invalid_tasks (taskid) VALUES (?)", (taskid,))

    def get_invalid_task(self, taskid: str) -> Optional[dict]:
        r = self._one("SELECT * FROM invalid_tasks WHERE taskid=?", (taskid,))
        return dict(r) if r else None

    # ---------------------------------------------------------------- task inputs
    def store_task_input(self, taskid: str, cid: str, data: Any) -> None:
        if self._one("SELECT taskid FROM task_inputs WHERE taskid=? AND cid=?", (taskid, cid)) is None:
            self._exec("INSERT INTO task_inputs (taskid, cid, data) VALUES (?,?,?)", (taskid, cid, json.dumps(data)))

    def get_task_input(self, taskid: str, cid: str) -> Optional[dict]:
        """db.ts:97-117 — the seed is (re)derived from the task id on every read."""
        r = self._one("SELECT data FROM task_inputs WHERE taskid=? AND cid=?", (taskid, cid))
        if r is None:
            return None
        data = json.loads(r["data"])
        data["seed"] = taskid2seed(taskid)
        return data

    # ---------------------------------------------------------------- solutions / contestations
    def store_solution(self, taskid: str, validator: str, blocktime: int, claimed: bool, cid: str) -> None:
        if self._one("SELECT id FROM solutions WHERE taskid=?", (taskid,)) is None:
            self._exec("INSERT INTO solutions (taskid, validator, blocktime, claimed, cid) VALUES (?,?,?,?,?)",
                       (taskid, validator, str(blocktime), claimed, cid))

    def get_solution(self, taskid: str) -> Optional[dict]:
        r = self._one("SELECT * FROM solutions WHERE taskid=?", (taskid,))
        return dict(r) if r else None

    def store_contestation(self, taskid: str, validator: str, blocktime: int, finish_start_index: int) -> None:
        if self._one("SELECT id FROM contestations WHERE taskid=?", (taskid,)) is None:
            self._exec("INSERT INTO contestations (taskid, validator, blocktime, finish_start_index) VALUES (?,?,?,?)",
                       (taskid, validator, str(blocktime), str(finish_start_index)))

    def get_contestation(self, taskid: str) -> Optional[dict]:
        r = self._one("SELECT * FROM contestations WHERE taskid=?", (taskid,))
        return dict(r) if r else None

    def store_contestation_vote(self, taskid: str, validator: str, yea: bool) -> None:
        if self._one("SELECT id FROM contestation_votes WHERE taskid=? AND validator=?", (taskid, validator)) is None:
            self._exec("INSERT INTO contestation_votes (taskid, validator, yea) VALUES (?,?,?)", (taskid, validator, yea))

    def get_contestation_votes(self, taskid: str) -> List[dict]:
        return [dict(r) for r in self._all("SELECT * FROM contestation_votes WHERE taskid=?", (taskid,))]

    # ---------------------------------------------------------------- models
    def store_model(self, mid: str, addr: str, mineable: bool, cid: str) -> None:
        self._exec("INSERT OR REPLACE INTO models (id, addr, mineable, cid) VALUES (?,?,?,?)", (mid, addr, mineable, cid))

    def get_model(self, mid: str) -> Optional[dict]:
        r = self._one("SELECT * FROM models WHERE id=?", (mid,))
        return dict(r) if r else None

    # ---------------------------------------------------------------- jobs
    def queue_job(self, method: str, priority: int = 0, waituntil: int = 0, concurrent: bool = False,
                  data: Optional[dict] = None) -> Job:
        """db.ts:237-267 dbQueueJob."""
        payload = json.dumps(data or {}, default=str)
        cur = self._exec("INSERT INTO jobs (priority, waituntil, concurrent, method, data) VALUES (?,?,?,?,?)",
                         (priority, waituntil, concurrent, method, payload))
        return Job(cur.lastrowid, priority, waituntil, concurrent, method, payload)

    def get_jobs(self, limit: int = 10000) -> List[Job]:
        """db.ts:131-144: ORDER BY priority DESC LIMIT ?"""
        rows = self._all("SELECT * FROM jobs ORDER BY priority DESC, id ASC LIMIT ?", (limit,))
        return [Job(r["id"], r["priority"], r["waituntil"], bool(r["concurrent"]), r["method"], r["data"]) for r in rows]

    def get_job(self, jobid: int) -> Optional[Job]:
        r = self._one("SELECT * FROM jobs WHERE id=?", (jobid,))
        return Job(r["id"], r["priority"], r["waituntil"], bool(r["concurrent"]), r["method"], r["data"]) if r else None

    def delete_job(self, jobid: int) -> None:
        self._exec("DELETE FROM jobs WHERE id=?", (jobid,))

    def clear_jobs_by_method(self, method: str) -> None:
        self._exec("DELETE FROM jobs WHERE method=?", (method,))

    def store_failed_job(self, job: Job) -> None:
        self._exec("INSERT INTO failed_jobs (method, data) VALUES (?,?)", (job.method, job.data))

    def get_failed_jobs(self) -> List[dict]:
        return [dict(r) for r in self._all("SELECT * FROM failed_jobs")]

    def requeue_failed_jobs(self, priority: int = 1) -> int:
        """Replay tool the reference lacks."""
        rows = self._all("SELECT * FROM failed_jobs")
        for r in rows:
            self._exec("INSERT INTO jobs (priority, waituntil, concurrent, method, data) VALUES (?,?,?,?,?)",
                       (priority, 0, False, r["method"], r["data"]))
        self._exec("DELETE FROM failed_jobs")
        return len(rows)
