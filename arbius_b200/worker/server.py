"""Cog-contract prediction server backed by the native models.

The unchanged reference miner posts ``{"input": {...hydrated fields..., "seed": int}}`` to the URL in
``ml.cog[modelid].url`` and expects ``{"output": ["data:<mime>;base64,..."]}`` with exactly one element
(reference: miner/src/index.ts:852-875). One server per GPU subset:

    python -m arbius_b200.worker.server --model anythingv3 --device cuda:0 --port 5000

Routes: POST /predictions (and POST /, some Cog builds use it), GET /health-check.
"""
from __future__ import annotations

import argparse
import json
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Dict

from ..miner.ml import encode_cog_output
from ..protocol.templates import load_template, output_mime

Predictor = Callable[[dict], bytes]


def handle_prediction(predict: Predictor, template: dict, body: dict) -> dict:
    inp = body.get("input")
    if not isinstance(inp, dict):
        raise ValueError("request body must be {\"input\": {...}}")
    if "seed" not in inp:
        raise ValueError("input.seed is required for reproducible outputs")   # docs/src/pages/register-model.mdx:23
    data = predict(inp)
    return encode_cog_output(data, output_mime(template))


def is_fatal_device_error(e: BaseException) -> bool:
    """A CUDA fault is sticky: the context is gone and every later prediction would fail the same way. The worker must
    die (non-zero) so its supervisor restarts it, instead of answering 422 for ever."""
    text = f"{type(e).__name__}: {e}".lower()
    if type(e).__name__ in ("AcceleratorError", "OutOfMemoryError"):
        return "out of memory" not in text           # OOM is recoverable; other accelerator errors are not
    return any(s in text for s in ("cuda error", "cudaerror", "device-side assert", "illegal memory access",
                                   "unspecified launch failure", "an illegal instruction", "launch failed",
                                   "misaligned address"))


def make_server(predict: Predictor, template: dict, host: str, port: int,
                on_fatal: Callable[[BaseException], None] = None) -> ThreadingHTTPServer:
    lock = threading.Lock()       # one prediction at a time per GPU subset
    stats = {"served": 0, "ms": 0.0}

    def die(e):
        import os
        import sys
        print(f"arbius_b200 worker: fatal device error, exiting for restart: {e}", file=sys.stderr, flush=True)
        os._exit(70)
    fatal = on_fatal or die

    class H(BaseHTTPRequestHandler):
        def log_message(self, *a):
            pass

        def _json(self, status, obj):
            data = json.dumps(obj).encode()
            self.send_response(status)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(data)))
            self.end_headers()
            self.wfile.write(data)

        def do_GET(self):
            if self.path in ("/health-check", "/"):
                return self._json(200, {"status": "READY", **stats})
            self._json(404, {"detail": "not found"})

        def do_POST(self):
            if self.path not in ("/predictions", "/"):
                return self._json(404, {"detail": "not found"})
            n = int(self.headers.get("Content-Length") or 0)
            try:
                body = json.loads(self.rfile.read(n) or b"{}")
                t0 = time.monotonic()
                with lock:
                    out = handle_prediction(predict, template, body)
                stats["served"] += 1
                stats["ms"] = (time.monotonic() - t0) * 1e3
                self._json(200, {"status": "succeeded", **out})
            except Exception as e:  # noqa: BLE001
                if is_fatal_device_error(e):
                    self._json(500, {"status": "failed", "detail": str(e), "fatal": True})
                    fatal(e)
                    return
                self._json(422, {"status": "failed", "detail": str(e)})

    return ThreadingHTTPServer((host, port), H)


def build_predictor(model: str, device: str, model_dir=None, tiny: bool = False) -> Predictor:
    from ..models.registry import build_predictor as _bp
    return _bp(model, device=device, model_dir=model_dir, tiny=tiny)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", required=True, help="template name: anythingv3 | kandinsky2 | zeroscopev2xl | damo | robust_video_matting")
    ap.add_argument("--device", default="cuda:0")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=5000)
    ap.add_argument("--model-dir", default=None)
    ap.add_argument("--tiny", action="store_true", help="tiny random-init config (tests)")
    a = ap.parse_args(argv)
    template = load_template(a.model)
    predict = build_predictor(a.model, a.device, a.model_dir, a.tiny)
    srv = make_server(predict, template, a.host, a.port)
    print(f"arbius_b200 worker: {a.model} on {a.device} listening on {a.host}:{a.port}", flush=True)
    srv.serve_forever()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
