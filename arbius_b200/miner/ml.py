"""ML strategy adapters: how a task's hydrated input becomes output files.

The reference has two strategies (reference: miner/src/ml.ts:6-10, miner/src/index.ts:752-779 Replicate,
:852-875 Cog): POST ``{"input": …}`` to a Cog container and decode ``output[0]`` (a data URI), or call the
Replicate API. Both are kept, byte-compatible, and a third is added:

  native   run the model in this process on the local B200s through ``arbius_b200.models`` (no HTTP,
           no base64 inflation, no fixed ``cache/out-1.png`` path — outputs are per-task byte strings).

Every backend returns ``{filename: bytes}`` keyed by the template's output filenames; nothing is
written to a shared cache directory, which is what made concurrent solves unsafe in the reference
(SURVEY §5.2).
"""
from __future__ import annotations

import base64
import json
import re
import threading
from typing import Callable, Dict, List, Optional

from ..protocol.hydrate import ModelEntry
from ..protocol.templates import output_filename

_DATA_URI = re.compile(r"^data:\w+/[\w.+-]+;base64,")


class MLError(Exception):
    pass


def decode_cog_output(res: dict) -> bytes:
    """index.ts:861-870: exactly one output element, a base64 data URI."""
    out = res.get("output") if isinstance(res, dict) else None
    if not isinstance(out, list) or len(out) != 1:
        raise MLError("unable to getfiles: expected exactly one output")
    payload = _DATA_URI.sub("", out[0])
    return base64.b64decode(payload)


def encode_cog_output(data: bytes, mime: str) -> dict:
    return {"output": [f"data:{mime};base64," + base64.b64encode(data).decode()]}


class MLBackend:
    name = "base"

    def getfiles(self, model: ModelEntry, taskid: str, inp: dict, endpoint: Optional[int] = None,
                 gpus: Optional[tuple] = None) -> Dict[str, bytes]:
        """``gpus`` is the subset the scheduler gave this solve (None: the backend picks an endpoint itself)."""
        raise NotImplementedError

    def endpoints(self, model: ModelEntry) -> int:
        """How many solves of this model can run at once."""
        return 1


class CogBackend(MLBackend):
    """``ml.cog[modelid] = {"url": …}`` as in the reference, plus an optional ``"urls": […]`` list
    (one Cog-contract worker per GPU subset) that the scheduler load-balances over."""
    name = "cog"

    def __init__(self, cog_cfg: dict, timeout: float = 120.0, post: Optional[Callable] = None):
        self.cfg = {k.lower(): v for k, v in cog_cfg.items()}
        self.timeout = timeout
        self._post = post or self._http_post
        self._rr: Dict[str, int] = {}
        self._lock = threading.Lock()

    @staticmethod
    def _http_post(url: str, payload: dict, timeout: float) -> dict:
        import urllib.request
        req = urllib.request.Request(url, data=json.dumps(payload).encode(), headers={"Content-Type": "application/json"})
        with urllib.request.urlopen(req, timeout=timeout) as r:
            return json.loads(r.read())

    def urls(self, model: ModelEntry) -> List[str]:
        ent = self.cfg.get(model.id.lower())
        if not ent:
            raise MLError(f"no cog url configured for model {model.id}")
        return list(ent.get("urls") or [ent["url"]])

    def endpoints(self, model: ModelEntry) -> int:
        return len(self.urls(model))

    def getfiles(self, model, taskid, inp, endpoint: Optional[int] = None, gpus=None):
        urls = self.urls(model)
        if endpoint is None and gpus:
            endpoint = gpus[0]                         # one worker per GPU: the scheduler's subset names it
        if endpoint is None:
            with self._lock:
                endpoint = self._rr.get(model.id, 0)
                self._rr[model.id] = (endpoint + 1) % len(urls)
        res = self._post(urls[endpoint % len(urls)], {"input": inp}, self.timeout)
        return {output_filename(model.template): decode_cog_output(res)}


class ReplicateBackend(MLBackend):
    """``replicate.run("ORG/TITLE:HASH", {input})`` derived from ``meta.docker`` (index.ts:763-771).
    Needs network + API token; present for configuration compatibility."""
    name = "replicate"

    def __init__(self, api_token: Optional[str], run: Optional[Callable] = None):
        self.api_token = api_token
        self._run = run

    @staticmethod
    def model_ref(template: dict) -> str:
        docker = template["meta"]["docker"]            # https://r8.im/org/title@sha256:hash
        m = re.search(r"r8\.im/([^@]+)@sha256:([0-9a-f]+)", docker)
        if not m:
            raise MLError(f"cannot derive replicate model from {docker}")
        return f"{m.group(1)}:{m.group(2)}"

    def getfiles(self, model, taskid, inp, endpoint=None, gpus=None):
        if self._run is None:
            raise MLError("replicate strategy needs network access and an api_token; use 'cog' or 'native'")
        out = self._run(self.model_ref(model.template), {"input": inp})
        data = out if isinstance(out, (bytes, bytearray)) else decode_cog_output({"output": [out] if isinstance(out, str) else out})
        return {output_filename(model.template): bytes(data)}


class _SlotLock:
    """A GPU's admission lock: up to ``slots`` single-GPU solves at once, or one exclusive holder (a subset solve)."""

    def __init__(self, slots: int = 1):
        self.slots, self._used = slots, 0
        self._cv = threading.Condition()

    def acquire_all(self) -> None:
        with self._cv:
            self._cv.wait_for(lambda: self._used == 0)
            self._used = self.slots

    def release_all(self) -> None:
        with self._cv:
            self._used = 0
            self._cv.notify_all()

    def locked(self) -> bool:
        return self._used > 0

    def __enter__(self):
        with self._cv:
            self._cv.wait_for(lambda: self._used < self.slots)
            self._used += 1
        return self

    def __exit__(self, *exc):
        with self._cv:
            self._used -= 1
            self._cv.notify_all()


class NativeBackend(MLBackend):
    """In-process predictors. ``predictors[modelid]`` is a list of callables ``inp -> {filename: bytes}``
    (one per GPU subset); each is guarded by its own lock so one GPU never runs two tasks at once."""
    name = "native"

    def __init__(self, predictors: Dict[str, List[Callable[[dict], Dict[str, bytes]]]],
                 subset_predictors: Optional[Dict[str, Dict[tuple, Callable]]] = None, slots: int = 1):
        """``subset_predictors[modelid][(g0, g1)]``: a predictor that spreads ONE task over that GPU subset (the
        guidance-parallel pair predictors of models/registry.py); used when the scheduler hands a solve that subset.
        ``slots`` > 1: that many solves may be inside one GPU's predictor at once (a ``BatchingPredictor`` merges them
        into one graph replay); a subset solve still takes its GPUs exclusively."""
        self.slots = max(1, int(slots))
        self.predictors = {k.lower(): list(v) for k, v in predictors.items()}
        self.subset_predictors = {k.lower(): {tuple(s): f for s, f in v.items()}
                                  for k, v in (subset_predictors or {}).items()}
        self._locks = {k: [_SlotLock(self.slots) for _ in v] for k, v in self.predictors.items()}
        self._rr: Dict[str, int] = {}
        self._lock = threading.Lock()

    def endpoints(self, model):
        return len(self.predictors.get(model.id.lower(), []))

    def subset_sizes(self, model) -> List[int]:
        """GPU counts one task of this model can be spread over (1 always; 2 when pair predictors were built)."""
        return sorted({1} | {len(s) for s in self.subset_predictors.get(model.id.lower(), {})})

    def getfiles(self, model, taskid, inp, endpoint: Optional[int] = None, gpus=None):
        mid = model.id.lower()
        preds = self.predictors.get(mid)
        if not preds:
            raise MLError(f"no native predictor for model {model.id}")
        if gpus and len(gpus) > 1:
            sub = self.subset_predictors.get(mid, {}).get(tuple(gpus))
            if sub is not None:
                locks = [self._locks[mid][g % len(preds)] for g in gpus]
                for lk in locks:                      # ascending GPU order everywhere: no lock cycles
                    lk.acquire_all()
                try:
                    out = sub(inp)
                except (MemoryError, ValueError) as e:
                    # the subset's peer heap is too small for this request, or its shape does not shard over the
                    # subset (models/video/pipeline.py UnsupportedSplit): one GPU does it
                    if isinstance(e, ValueError) and type(e).__name__ != "UnsupportedSplit":
                        raise
                    out = None
                finally:
                    for lk in reversed(locks):
                        lk.release_all()
                if isinstance(out, (bytes, bytearray)):
                    out = {output_filename(model.template): bytes(out)}
                if out is not None:
                    return out
        if endpoint is None and gpus:
            endpoint = gpus[0]
        if endpoint is None:
            with self._lock:
                endpoint = self._rr.get(model.id, 0)
                self._rr[model.id] = (endpoint + 1) % len(preds)
        i = endpoint % len(preds)
        with self._locks[mid][i]:
            out = preds[i](inp)
        if isinstance(out, (bytes, bytearray)):
            out = {output_filename(model.template): bytes(out)}
        return out


def make_backend(c, native_predictors: Optional[dict] = None, subset_predictors: Optional[dict] = None,
                 slots: int = 1) -> MLBackend:
    s = c.ml.strategy
    if s == "cog":
        return CogBackend(c.ml.cog)
    if s == "replicate":
        return ReplicateBackend((c.ml.replicate or {}).get("api_token"))
    if s == "native":
        return NativeBackend(native_predictors or {}, subset_predictors, slots)
    raise MLError(f"unknown ml strategy {s!r}")
