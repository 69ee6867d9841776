"""Robust Video Matting pipeline: ``input_video`` + ``output_type`` -> ``out-1.mp4`` (templates/robust_video_matting.json).

The network is recurrent, so frames are processed in order; one frame = one CUDA-graph replay (static u8 input and
output buffers, the four ConvGRU states updated in place). The whole clip is uploaded once from pinned memory, the
composited frames are read back once.
"""
from __future__ import annotations

import io
import os
import tempfile
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ...ops.matting import MODES
from .ref import MattingNetwork, RvmConfig, init_rvm_

__all__ = ["RvmConfig", "RvmPipeline", "decode_video", "composite_reference"]


def decode_video(data) -> Tuple[np.ndarray, int]:
    """bytes / path / ndarray -> (frames [T,H,W,3] uint8 RGB, fps). MJPEG-in-MP4 (what this framework writes) is decoded
    with the in-tree demuxer + PIL; anything else goes through OpenCV's container reader when it is importable."""
    if isinstance(data, np.ndarray):
        assert data.ndim == 4 and data.shape[-1] == 3 and data.dtype == np.uint8
        return data, 30
    if isinstance(data, (str, os.PathLike)):
        with open(data, "rb") as f:
            data = f.read()
    from ...codecs.mp4 import demux_samples, probe_fps
    try:
        samples = demux_samples(data)
    except Exception:
        samples = []
    if samples and samples[0][:2] == b"\xff\xd8":
        from PIL import Image
        frames = [np.asarray(Image.open(io.BytesIO(s)).convert("RGB")) for s in samples]
        return np.stack(frames), probe_fps(data)
    try:
        import cv2
    except ImportError as e:                                    # pragma: no cover
        raise ValueError("input_video is not MJPEG-in-MP4 and no container decoder (cv2) is available") from e
    with tempfile.NamedTemporaryFile(suffix=".mp4") as tmp:
        tmp.write(data)
        tmp.flush()
        cap = cv2.VideoCapture(tmp.name)
        fps = int(round(cap.get(cv2.CAP_PROP_FPS) or 30)) or 30
        frames = []
        while True:
            ok, fr = cap.read()
            if not ok:
                break
            frames.append(fr[:, :, ::-1])
        cap.release()
    if not frames:
        raise ValueError("input_video could not be decoded")
    return np.ascontiguousarray(np.stack(frames)), fps


def composite_reference(fgr: torch.Tensor, pha: torch.Tensor, mode: int) -> torch.Tensor:
    """fgr [N,3,H,W], pha [N,1,H,W] in [0,1] -> uint8 [N,H,W,3]; the eager twin of the native compositing stage."""
    bg = torch.tensor([120.0, 255.0, 155.0], device=fgr.device, dtype=fgr.dtype).view(1, 3, 1, 1) / 255.0
    if mode == 0:
        v = fgr * pha + bg * (1 - pha)
    elif mode == 1:
        v = pha.expand(-1, 3, -1, -1)
    else:
        v = fgr * pha
    return torch.round(v.float() * 255.0).clamp(0, 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous()


class RvmPipeline:
    def __init__(self, cfg: RvmConfig = RvmConfig(), backend: str = "native", device: str = "cuda", dtype=torch.float16,
                 seed: int = 0, model_dir: Optional[str] = None, use_graph: bool = True):
        self.cfg, self.backend, self.device, self.dtype = cfg, backend, torch.device(device), dtype
        net = MattingNetwork(cfg)
        if model_dir:
            path = os.path.join(model_dir, "rvm_mobilenetv3.pth")
            net.load_state_dict(torch.load(path, map_location="cpu", weights_only=True), strict=True)
            net.eval().requires_grad_(False)
        else:
            init_rvm_(net, seed + 40)
        if backend == "eager":
            self.net = net.to(device=self.device, dtype=dtype)
        else:
            from .native import NativeRvm
            self.net = NativeRvm(net.state_dict(), cfg, dtype, device)
        self.use_graph = use_graph and backend == "native"
        self.chunk_frames = 8            # frames per host<->device staging chunk (two chunks in flight each way)
        self._graphs: Dict[Tuple, dict] = {}
        self.h2d_bytes = self.d2h_bytes = 0
        self.last_launches = 0

    # ------------------------------------------------------------------------------------------------
    def matte(self, frames: np.ndarray, output_type: str = "green-screen") -> np.ndarray:
        """frames [T,H,W,3] uint8 -> composited frames [T,H,W,3] uint8"""
        mode = MODES[output_type or "green-screen"]
        if self.backend == "eager":
            return self._eager(frames, mode)
        return self._native(frames, mode)

    def _eager(self, frames: np.ndarray, mode: int) -> np.ndarray:
        rec: List[Optional[torch.Tensor]] = [None] * 4
        out = []
        x = torch.from_numpy(frames).to(self.device)
        for t in range(x.shape[0]):
            src = (x[t:t + 1].permute(0, 3, 1, 2).to(self.dtype) / 255.0)
            fgr, pha, *rec = self.net(src, *rec)
            out.append(composite_reference(fgr, pha, mode))
        return torch.cat(out).cpu().numpy()

    def _state_for(self, T: int, H: int, W: int, mode: int) -> dict:
        key = (H, W, mode)
        st = self._graphs.get(key)
        if st is None:
            st = {"src": torch.zeros((1, H, W, 3), dtype=torch.uint8, device=self.device),
                  "out": torch.zeros((1, H, W, 3), dtype=torch.uint8, device=self.device),
                  "rec": self.net.new_state(1, H, W), "graph": None}
            self._graphs[key] = st
        return st

    def _frame(self, st: dict, mode: int) -> None:
        self.net.forward(st["src"], st["rec"], mode=mode, want_fgr_pha=False, out_u8=st["out"])

    def _native(self, frames: np.ndarray, mode: int) -> np.ndarray:
        from ...ops import native as N
        T, H, W, _ = frames.shape
        st = self._state_for(T, H, W, mode)
        frames = np.ascontiguousarray(frames)
        self.h2d_bytes = self.d2h_bytes = frames.size
        CH = self.chunk_frames
        if "pin_in" not in st:          # two pinned + two device staging chunks each way, reused by every clip of this size
            st["pin_in"] = [torch.empty((CH, H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
            st["pin_out"] = [torch.empty((CH, H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
            st["dev_in"] = [torch.empty((CH, H, W, 3), dtype=torch.uint8, device=self.device) for _ in range(2)]
            st["dev_out"] = [torch.empty((CH, H, W, 3), dtype=torch.uint8, device=self.device) for _ in range(2)]
            st["copy_stream"] = torch.cuda.Stream(device=self.device)
        for r in st["rec"]:
            r.zero_()
        lib = N.load()
        c0 = lib.ab_launch_count() if hasattr(lib, "ab_launch_count") else 0
        if self.use_graph and st["graph"] is None:
            st["src"].copy_(torch.from_numpy(frames[0:1]).to(self.device))
            s = torch.cuda.Stream(device=self.device)
            s.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(s):
                self._frame(st, mode)                      # warm-up (allocations, lazy loads) outside capture
            torch.cuda.current_stream(self.device).wait_stream(s)
            torch.cuda.synchronize(self.device)
            c1 = lib.ab_launch_count() if hasattr(lib, "ab_launch_count") else 0
            st["per_frame"] = c1 - c0
            for r in st["rec"]:
                r.zero_()
            g = torch.cuda.CUDAGraph()
            with N.graph_capture(g):
                self._frame(st, mode)
            st["graph"] = g
            for r in st["rec"]:
                r.zero_()
        # Chunked pipeline: while the GPU mattes chunk k (compute stream), the copy stream brings chunk k+1 in and takes
        # chunk k-1 out, and this thread fills / drains the pinned staging buffers. The recurrence only orders frames
        # on the compute stream.
        main, cs = torch.cuda.current_stream(self.device), st["copy_stream"]
        result = np.empty_like(frames)
        nchunk = (T + CH - 1) // CH
        h2d_done = [torch.cuda.Event() for _ in range(2)]
        comp_done = [torch.cuda.Event() for _ in range(2)]
        d2h_done = [torch.cuda.Event() for _ in range(2)]

        def stage_in(k: int) -> None:
            b, lo, hi = k % 2, k * CH, min(T, (k + 1) * CH)
            if k >= 2:
                comp_done[b].synchronize()                  # dev_in[b] / pin_in[b] were last used by chunk k-2
            np.copyto(st["pin_in"][b].numpy()[:hi - lo], frames[lo:hi])
            with torch.cuda.stream(cs):
                st["dev_in"][b][:hi - lo].copy_(st["pin_in"][b][:hi - lo], non_blocking=True)
                h2d_done[b].record(cs)

        def drain(k: int) -> None:
            b, lo, hi = k % 2, k * CH, min(T, (k + 1) * CH)
            d2h_done[b].synchronize()
            np.copyto(result[lo:hi], st["pin_out"][b].numpy()[:hi - lo])

        stage_in(0)
        for k in range(nchunk):
            b, lo, hi = k % 2, k * CH, min(T, (k + 1) * CH)
            if k + 1 < nchunk:
                stage_in(k + 1)
            main.wait_event(h2d_done[b])
            if k >= 2:
                main.wait_event(d2h_done[b])                # dev_out[b] still being read out for chunk k-2
            for t in range(hi - lo):
                st["src"].copy_(st["dev_in"][b][t:t + 1])
                if st["graph"] is not None:
                    st["graph"].replay()
                else:
                    self._frame(st, mode)
                st["dev_out"][b][t:t + 1].copy_(st["out"])
            comp_done[b].record(main)
            with torch.cuda.stream(cs):
                cs.wait_event(comp_done[b])
                st["pin_out"][b][:hi - lo].copy_(st["dev_out"][b][:hi - lo], non_blocking=True)
                d2h_done[b].record(cs)
            if k >= 1:
                drain(k - 1)
        drain(nchunk - 1)
        self.last_launches = st.get("per_frame", 0) * T
        return result

    # ------------------------------------------------------------------------------------------------
    def predict(self, inp: dict) -> bytes:
        """Hydrated template input -> bytes of ``out-1.mp4``."""
        from ...codecs.mp4 import encode_mp4
        frames, fps = decode_video(inp["input_video"])
        out = self.matte(frames, inp.get("output_type") or "green-screen")
        return encode_mp4(out, int(inp.get("fps", fps)))
