"""Operator HTTP RPC — same routes and JSON shapes as the reference (reference: miner/src/rpc.ts:11-104).

  GET  /                   "Arbius Miner RPC"
  POST /api/jobs/queue     typed field check, then dbQueueJob
  POST /api/jobs/get       {limit?}  (the docs call it /api/jobs/list — both are served)
  POST /api/jobs/delete    {id}
  POST /api/db/run         unvalidated queue (kept for compatibility)
  GET  /metrics            additive: Prometheus text (solved tasks, latency, queue depth)
stdlib ThreadingHTTPServer: no dependency beyond Python.
"""
from __future__ import annotations

import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Optional

from .db import MinerDB

_JS_TYPES = {"string": str, "number": (int, float), "boolean": bool, "object": (dict, list)}


def _js_typeof(v) -> str:
    if isinstance(v, bool):
        return "boolean"
    if isinstance(v, (int, float)):
        return "number"
    if isinstance(v, str):
        return "string"
    if v is None:
        return "object"
    return "object"


def handle(db: MinerDB, method: str, path: str, body: dict, node=None):
    """Pure request handler (unit-testable without sockets). Returns (status, payload, content_type)."""
    if method == "GET" and path == "/":
        return 200, "Arbius Miner RPC", "text/html"
    if method == "GET" and path == "/metrics":
        lines = [f"arbius_jobs_queued {len(db.get_jobs())}", f"arbius_jobs_failed {len(db.get_failed_jobs())}"]
        if node is not None:
            lat = sorted(s["total_ms"] for s in node.solved)
            lines.append(f"arbius_tasks_solved_total {len(lat)}")
            if lat:
                from ..utils import LatencyHistogram
                lines.append(f"arbius_task_latency_ms_p50 {lat[len(lat) // 2]:.3f}")
                hist = LatencyHistogram()
                for ms in lat:
                    hist.observe(ms)
                lines.append(hist.prometheus("arbius_task_latency_ms"))
            lines.append(f"arbius_gpu_tasks_inflight {node.pool.inflight}")
        return 200, "\n".join(lines) + "\n", "text/plain"
    if method != "POST":
        return 404, {"status": "fail", "e": "not found"}, "application/json"
    try:
        if path == "/api/jobs/queue":
            for fld, etype in (("method", "string"), ("priority", "number"), ("waituntil", "number"),
                               ("concurrent", "boolean"), ("data", "object")):
                if fld not in body:
                    raise ValueError(f"missing required field ({fld})")
                if _js_typeof(body[fld]) != etype:
                    raise ValueError(f"expected ({fld}) to be type ({etype}), got ({_js_typeof(body[fld])})")
            db.queue_job(body["method"], int(body["priority"]), int(body["waituntil"]), body["concurrent"], body["data"])
            return 200, {"status": "ok"}, "application/json"
        if path in ("/api/jobs/get", "/api/jobs/list"):
            limit = 100_000
            if "limit" in body:
                try:
                    limit = int(body["limit"])
                except (TypeError, ValueError):
                    raise ValueError("limit NaN")
            jobs = [j.__dict__ for j in db.get_jobs(limit)]
            return 200, {"status": "ok", "jobs": jobs}, "application/json"
        if path == "/api/jobs/delete":
            try:
                jid = int(body.get("id"))
            except (TypeError, ValueError):
                raise ValueError("id NaN")
            db.delete_job(jid)
            return 200, {"status": "ok"}, "application/json"
        if path == "/api/db/run":
            db.queue_job(body.get("method"), body.get("priority") or 0, body.get("waituntil") or 0,
                         bool(body.get("concurrent")), body.get("data"))
            return 200, {"status": "ok"}, "application/json"
    except Exception as e:  # noqa: BLE001 - the reference answers 200 {status: fail}
        return 200, {"status": "fail", "e": f"Error: {e}"}, "application/json"
    return 404, {"status": "fail", "e": "not found"}, "application/json"


def make_server(db: MinerDB, host: str, port: int, node=None) -> ThreadingHTTPServer:
    class H(BaseHTTPRequestHandler):
        def log_message(self, *a):  # quiet
            pass

        def _reply(self, status, payload, ctype):
            data = payload.encode() if isinstance(payload, str) else json.dumps(payload).encode()
            self.send_response(status)
            self.send_header("Content-Type", ctype)
            self.send_header("Content-Length", str(len(data)))
            self.end_headers()
            self.wfile.write(data)

        def do_GET(self):
            self._reply(*handle(db, "GET", self.path, {}, node))

        def do_POST(self):
            n = int(self.headers.get("Content-Length") or 0)
            try:
                body = json.loads(self.rfile.read(n) or b"{}")
            except json.JSONDecodeError:
                body = {}
            self._reply(*handle(db, "POST", self.path, body if isinstance(body, dict) else {}, node))

    return ThreadingHTTPServer((host, port), H)


def initialize_rpc(db: MinerDB, host: str, port: int, node=None) -> ThreadingHTTPServer:
    srv = make_server(db, host, port, node)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    return srv
