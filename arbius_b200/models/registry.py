"""Template name -> predictor factory. A predictor maps the hydrated template input (+ ``seed``) to the
bytes of the template's output file (``out-1.png`` / ``out-1.mp4``); SURVEY Appendix A.3.

One process may own several GPUs (the miner's GpuPool runs one predictor per device on worker threads), and the native
ops launch on the *current* device's stream: pipelines are therefore built, and every prediction runs, under a
``torch.cuda.device`` guard for the predictor's own device.
"""
from __future__ import annotations

from typing import Callable, Optional

FAMILIES = ("anythingv3", "kandinsky2", "zeroscopev2xl", "damo", "robust_video_matting")


def _make_pipeline(name: str, device: str, model_dir: Optional[str], tiny: bool, backend: str, dtype):
    if name == "anythingv3":
        from .sd15.pipeline import SD15Config, SD15Pipeline
        return SD15Pipeline(SD15Config.tiny() if tiny else SD15Config(), backend=backend, device=device, dtype=dtype,
                            model_dir=model_dir)
    if name == "kandinsky2":
        from .kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
        return Kandinsky2Pipeline(Kandinsky2Config.tiny() if tiny else Kandinsky2Config(), backend=backend,
                                  device=device, dtype=dtype, model_dir=model_dir)
    if name in ("zeroscopev2xl", "damo"):
        from .video.pipeline import TextToVideoConfig, TextToVideoPipeline
        cfg = TextToVideoConfig.tiny() if tiny else (TextToVideoConfig.zeroscope_xl() if name == "zeroscopev2xl"
                                                      else TextToVideoConfig.damo())
        return TextToVideoPipeline(cfg, backend=backend, device=device, dtype=dtype, model_dir=model_dir, family=name)
    if name == "robust_video_matting":
        from .rvm.pipeline import RvmConfig, RvmPipeline
        return RvmPipeline(RvmConfig.tiny() if tiny else RvmConfig(), backend=backend, device=device, dtype=dtype,
                           model_dir=model_dir)
    raise ValueError(f"unknown model family {name!r}; choices: {FAMILIES}")


def build_predictor(name: str, device: str = "cuda:0", model_dir: Optional[str] = None, tiny: bool = False,
                    backend: Optional[str] = None, dtype=None, task_batch: int = 1) -> Callable[[dict], bytes]:
    """``task_batch`` > 1 (families in ``BATCH_FAMILIES``, native backend) returns a ``BatchingPredictor``: a callable
    that may be entered by ``task_batch`` threads at once and runs their tasks in one graph replay."""
    import torch
    dev = torch.device(device)
    backend = backend or ("native" if dev.type == "cuda" else "eager")
    dtype = dtype or (torch.float16 if dev.type == "cuda" else torch.float32)
    if dev.type != "cuda":
        return _make_pipeline(name, device, model_dir, tiny, backend, dtype).predict
    with torch.cuda.device(dev):
        pipe = _make_pipeline(name, device, model_dir, tiny, backend, dtype)
    if task_batch > 1 and backend == "native" and name in BATCH_FAMILIES:
        return BatchingPredictor(pipe, dev, task_batch)

    import threading
    serial = threading.Lock()          # a pipeline's buffers and graphs serve one task at a time

    def predict(inp: dict) -> bytes:
        with serial, torch.cuda.device(dev):
            return pipe.predict(inp)

    predict.pipeline = pipe           # the miner's cost model and the tests reach the pipeline through the callable
    return predict


class BatchingPredictor:
    """Dynamic task batching in front of a pipeline that has ``generate_batch`` (the SD-1.5 family): callers block in
    ``__call__``; one worker thread per GPU takes the oldest request, adds up to ``task_batch - 1`` queued requests with
    the same (size, steps, guidance, sampler) — waiting at most ``window_ms`` for them when more solves are known to be
    on their way — and runs the group as ONE graph replay. Every task's bytes are the ones it gets alone, so batching
    is invisible on chain; tasks/hour go up because the low-resolution UNet levels are launch-latency bound."""

    def __init__(self, pipe, device=None, task_batch: int = 2, window_ms: float = 2.0, encode=None):
        import queue
        import threading
        self.pipeline, self.device = pipe, device
        self.task_batch, self.window_s = max(1, int(task_batch)), window_ms / 1e3
        self._q: "queue.Queue" = queue.Queue()
        self._encode = encode
        self.batches = []                                  # sizes of the groups that ran (metrics, tests)
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()

    def __call__(self, inp: dict) -> bytes:
        from concurrent.futures import Future
        fut: Future = Future()
        self._q.put((self.pipeline._task_args(inp), fut))
        return fut.result()

    @staticmethod
    def _key(t: dict):
        return (t["width"], t["height"], t["num_inference_steps"], t["guidance_scale"], t["scheduler"])

    def _loop(self) -> None:
        import contextlib
        import queue
        import time
        held = []                                          # requests taken off the queue that did not fit a group
        while True:
            first = held.pop(0) if held else self._q.get()
            if first is None:
                return
            group, rest, deadline, stop = [first], [], None, False
            while len(group) < self.task_batch:
                if held:
                    nxt = held.pop(0)
                else:
                    deadline = deadline or (time.monotonic() + self.window_s)
                    try:
                        nxt = self._q.get(timeout=max(0.0, deadline - time.monotonic()))
                    except queue.Empty:
                        break
                if nxt is None:
                    stop = True
                    break
                (group if self._key(nxt[0]) == self._key(first[0]) else rest).append(nxt)
            n = 1 << (len(group).bit_length() - 1)         # group sizes are powers of two: one graph per size in use
            held = group[n:] + rest + held
            group = group[:n]
            self.batches.append(len(group))
            try:
                ctx = contextlib.nullcontext()
                if self.device is not None:
                    import torch
                    ctx = torch.cuda.device(self.device)
                with ctx:
                    imgs = self.pipeline.generate_batch([t for t, _ in group])
                if self._encode is None:
                    from ..codecs.png import encode_png
                    self._encode = encode_png
                for (_, fut), img in zip(group, imgs):
                    fut.set_result(self._encode(img))
            except BaseException as e:                     # noqa: BLE001 - delivered to every waiting caller
                for _, fut in group:
                    if not fut.done():
                        fut.set_exception(e)
            if stop:
                held.append(None)

    def close(self) -> None:
        self._q.put(None)
        self._thread.join(5)


BATCH_FAMILIES = ("anythingv3",)
PAIR_FAMILIES = ("anythingv3", "kandinsky2", "zeroscopev2xl", "damo")


def build_pair_predictor(name: str, devices=(0, 1), model_dir: Optional[str] = None, tiny: bool = False,
                         dtype=None) -> Callable[[dict], bytes]:
    """One task over TWO GPUs of this process (a 2-GPU subset of the miner's pool): guidance parallelism — each device
    runs one classifier-free-guidance branch per UNet step, predictions cross through peer memory, both produce the
    same bytes as the single-GPU pipeline; the first device's output is returned. Lower latency, not more throughput:
    the scheduler only picks it when GPUs would otherwise idle."""
    import threading

    import torch

    from ..parallel.symm import PeerHeap
    if name not in PAIR_FAMILIES:
        raise ValueError(f"{name!r} has no 2-GPU form; choices: {PAIR_FAMILIES}")
    assert len(devices) == 2
    dtype = dtype or torch.float16
    heaps = PeerHeap.local_group(64 << 20, devices)
    pipes = []
    for d, heap in zip(devices, heaps):
        with torch.cuda.device(d):
            if name == "anythingv3":
                from .sd15.pipeline import SD15Config, SD15Pipeline
                pipes.append(SD15Pipeline(SD15Config.tiny() if tiny else SD15Config(), backend="native", device=f"cuda:{d}",
                                          dtype=dtype, model_dir=model_dir, cfg_parallel=heap))
            elif name == "kandinsky2":
                from .kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
                pipes.append(Kandinsky2Pipeline(Kandinsky2Config.tiny() if tiny else Kandinsky2Config(), backend="native",
                                                device=f"cuda:{d}", dtype=dtype, model_dir=model_dir, cfg_parallel=heap))
            else:
                from .video.pipeline import TextToVideoConfig, TextToVideoPipeline
                cfg = TextToVideoConfig.tiny() if tiny else (TextToVideoConfig.zeroscope_xl() if name == "zeroscopev2xl"
                                                              else TextToVideoConfig.damo())
                pipes.append(TextToVideoPipeline(cfg, backend="native", device=f"cuda:{d}", dtype=dtype,
                                                 model_dir=model_dir, family=name, cfg_parallel=heap))

    def predict(inp: dict) -> bytes:
        out = [None, None]

        def run(i: int) -> None:
            try:
                with torch.cuda.device(devices[i]):
                    out[i] = pipes[i].predict(dict(inp))
            except BaseException as e:                      # noqa: BLE001 - re-raised on the calling thread
                import time
                e._when = time.monotonic()
                out[i] = e
        threads = [threading.Thread(target=run, args=(i,)) for i in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        errs = sorted((o for o in out if isinstance(o, BaseException)), key=lambda e: getattr(e, "_when", 0.0))
        if errs:                                            # the first failure is the cause; a peer's barrier timeout follows it
            raise errs[0]
        return out[0]

    predict.pipelines = pipes
    return predict


FRAME_FAMILIES = ("zeroscopev2xl", "damo")


def build_frame_parallel_predictor(name: str, devices=(0, 1), model_dir: Optional[str] = None, tiny: bool = False,
                                   dtype=None, heap_bytes: Optional[int] = None) -> Callable[[dict], bytes]:
    """One text-to-video task over 2, 4 or 8 GPUs of this process: sequence (frame) parallelism — each device denoises
    and decodes F/P frames, the temporal layers cross devices through peer memory (parallel/frame_shard.py), and the
    clip has the bytes of the single-GPU pipeline. The first device assembles and returns the MP4."""
    import threading

    import torch

    from ..parallel.symm import PeerHeap
    from .video.pipeline import TextToVideoConfig, TextToVideoPipeline
    if name not in FRAME_FAMILIES:
        raise ValueError(f"{name!r} has no frame-parallel form; choices: {FRAME_FAMILIES}")
    assert len(devices) in (2, 4, 8)
    dtype = dtype or torch.float16
    # sized for the template's largest default clip (ZeroScope XL 1024x576x24 needs 1.4 GB); a request that does not
    # fit raises MemoryError on every rank before the first barrier, and the miner falls back to one GPU
    heaps = PeerHeap.local_group(heap_bytes or ((1 << 30) if tiny else (3 << 30)), devices)
    cfg = TextToVideoConfig.tiny() if tiny else (TextToVideoConfig.zeroscope_xl() if name == "zeroscopev2xl"
                                                  else TextToVideoConfig.damo())
    pipes = []
    for d, heap in zip(devices, heaps):
        with torch.cuda.device(d):
            pipes.append(TextToVideoPipeline(cfg, backend="native", device=f"cuda:{d}", dtype=dtype, model_dir=model_dir,
                                             family=name, frame_parallel=heap))

    def predict(inp: dict) -> bytes:
        out = [None] * len(devices)

        def run(i: int) -> None:
            try:
                with torch.cuda.device(devices[i]):
                    out[i] = pipes[i].predict(dict(inp))
            except BaseException as e:                      # noqa: BLE001 - re-raised on the calling thread
                import time
                e._when = time.monotonic()
                out[i] = e
        threads = [threading.Thread(target=run, args=(i,)) for i in range(len(devices))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        errs = sorted((o for o in out if isinstance(o, BaseException)), key=lambda e: getattr(e, "_when", 0.0))
        if errs:                                            # the first failure is the cause; a peer's barrier timeout follows it
            raise errs[0]
        return out[0]

    predict.pipelines = pipes
    return predict
