"""Text-to-video pipeline for ZeroScope v2 XL and DAMO text-to-video-ms-1.7b: prompt -> frames -> MP4 bytes.

Stands in for the ``anotherjesse/zeroscope-v2-xl`` and ``cjwbw/damo-text-to-video`` Cog containers (reference:
templates/zeroscopev2xl.json, templates/damo.json; request/response contract miner/src/index.ts:852-875):
OpenCLIP ViT-H text encoder -> N UNet3D steps at batch 2 (CFG) -> sampler -> per-frame AutoencoderKL decode ->
uint8 frames -> deterministic MP4 (``codecs/mp4.py``). Same two backends, RNG and sampler plans as the image
pipeline; latents are channels-last [1, F, h, w, 4] so frame shards and row bands are contiguous index ranges.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ... import ops
from ..schedulers import make_scheduler, apply_plan_torch
from ..sd15.clip_ref import CLIPTextModel, ClipTextConfig
from ..sd15.unet_ref import init_weights_
from ..sd15.vae_ref import AutoencoderKLDecoder, VaeConfig
from ..tokenizer import load_tokenizer
from .unet3d_ref import UNet3DConditionModel, UNet3DConfig


@dataclass
class TextToVideoConfig:
    unet: UNet3DConfig = field(default_factory=UNet3DConfig)
    clip: ClipTextConfig = field(default_factory=ClipTextConfig.openclip_h)
    vae: VaeConfig = field(default_factory=VaeConfig)
    default_scheduler: str = "DPMSolverMultistep"
    default_guidance: float = 17.5
    default_size: Tuple[int, int] = (1024, 576)
    default_frames: int = 24
    default_fps: int = 24

    @staticmethod
    def zeroscope_xl() -> "TextToVideoConfig":
        return TextToVideoConfig()

    @staticmethod
    def damo() -> "TextToVideoConfig":
        return TextToVideoConfig(default_scheduler="DDIM", default_guidance=9.0, default_size=(256, 256),
                                 default_frames=16, default_fps=8)

    @staticmethod
    def tiny() -> "TextToVideoConfig":
        return TextToVideoConfig(UNet3DConfig.tiny(), ClipTextConfig.tiny(), VaeConfig.tiny(), default_size=(32, 32),
                                 default_frames=8, default_fps=8)


class UnsupportedSplit(ValueError):
    """This task's shape cannot be spread over the ranks it was given (the caller should use one GPU)."""


def frame_split_ok(nf: int, h: int, w: int, ranks: int, levels: int = 4) -> bool:
    """Frames must divide over the ranks, 2..128 frames fit one temporal-attention tile, and at every UNet level the
    rows a rank holds (frames/ranks x pixels) must split evenly into its 8/ranks canonical GroupNorm shards."""
    if ranks not in (2, 4, 8) or nf % ranks or not 2 <= nf <= 128:
        return False
    per = 8 // ranks
    for _ in range(levels):
        if ((nf // ranks) * h * w) % per:
            return False
        h, w = (h + 1) // 2, (w + 1) // 2
    return True


class TextToVideoPipeline:
    def __init__(self, cfg: TextToVideoConfig, backend: str = "native", device: str = "cuda", dtype=torch.float16,
                 seed: int = 0, model_dir: Optional[str] = None, family: str = "zeroscopev2xl", use_graph: bool = True,
                 cfg_parallel=None, frame_parallel=None):
        """``cfg_parallel``: a 2-rank ``PeerHeap`` — guidance parallelism as in ``SD15Pipeline``: each GPU runs one
        classifier-free-guidance branch of every UNet3D step and the predictions are exchanged through peer memory.
        The video kernels are large, so one task takes close to half the single-GPU time, with identical bytes.
        ``frame_parallel``: a 2/4/8-rank ``PeerHeap`` — sequence parallelism (parallel/frame_shard.py): every rank
        denoises and decodes F/P frames of ONE task; temporal GroupNorm, temporal conv and temporal attention cross
        ranks through peer memory in the single-GPU kernels' canonical order, so the bytes are those of one GPU.
        Rank 0 returns the clip, the other ranks return None."""
        self.cfg, self.backend, self.device, self.dtype, self.family = cfg, backend, torch.device(device), dtype, family
        self.cfgp = cfg_parallel
        self.fp = frame_parallel
        self._shard = None
        assert cfg_parallel is None or (cfg_parallel.world == 2 and backend == "native"), "cfg_parallel needs a 2-rank heap"
        assert frame_parallel is None or backend == "native", "frame_parallel needs the native backend"
        # both may be given (two heaps over different rank groups): guidance branch x frame shard, 2*P ranks per task
        self.tokenizer = load_tokenizer(model_dir, cfg.clip.vocab_size)
        with torch.device(device):
            unet, clip, vae = UNet3DConditionModel(cfg.unet), CLIPTextModel(cfg.clip), AutoencoderKLDecoder(cfg.vae)
        if model_dir:
            from pathlib import Path

            from safetensors.torch import load_file
            for mod, sub in ((unet, "unet"), (clip, "text_encoder"), (vae, "vae")):
                sd = {}
                for fpath in sorted((Path(model_dir) / sub).glob("*.safetensors")):
                    sd.update(load_file(str(fpath)))
                mod.load_state_dict({k: sd[k] for k in mod.state_dict()}, strict=True)
        else:
            for i, m in enumerate((unet, clip, vae)):
                init_weights_(m, seed + 10 + i)
        for m in (unet, clip, vae):
            m.to(dtype).eval().requires_grad_(False)
        self.vae_scale = 2 ** (len(cfg.vae.block_out_channels) - 1)
        self.use_graph = use_graph and backend == "native"
        self._graphs: Dict[Tuple, dict] = {}
        if backend == "eager":
            self.unet, self.clip, self.vae = unet, clip, vae
        else:
            from ..sd15.native import NativeClipText, NativeVaeDecoder
            from .native import NativeUNet3D
            self.unet = NativeUNet3D(unet.state_dict(), cfg.unet, dtype, device)
            self.clip = NativeClipText(clip.state_dict(), cfg.clip, dtype, device)
            self.vae = NativeVaeDecoder(vae.state_dict(), cfg.vae, dtype, device)
            del unet, clip, vae
        self.h2d_bytes = self.d2h_bytes = 0
        self.last_launches = 0

    def _ids(self, prompt: str, negative: str) -> torch.Tensor:
        return torch.tensor([self.tokenizer.encode(negative), self.tokenizer.encode(prompt)], dtype=torch.int64)

    # ------------------------------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, prompt: str, negative_prompt: str = "", num_frames: Optional[int] = None,
                 num_inference_steps: int = 50, width: Optional[int] = None, height: Optional[int] = None,
                 guidance_scale: Optional[float] = None, scheduler: Optional[str] = None, seed: int = 0) -> np.ndarray:
        c = self.cfg
        nf = int(num_frames or c.default_frames)
        width, height = int(width or c.default_size[0]), int(height or c.default_size[1])
        g = float(guidance_scale if guidance_scale is not None else c.default_guidance)
        sched = scheduler or c.default_scheduler
        ids = self._ids(prompt, negative_prompt)
        if self.backend == "eager":
            return self._eager(ids, nf, num_inference_steps, width, height, g, sched, seed)
        return self._native(ids, nf, num_inference_steps, width, height, g, sched, seed)

    def _eager(self, ids, nf, steps, width, height, g, sched, seed) -> np.ndarray:
        dev, dt = self.device, self.dtype
        h, w = height // self.vae_scale, width // self.vae_scale
        sch = make_scheduler(sched)
        sch.set_timesteps(steps)
        ctx = self.clip(ids.to(dev))
        self.h2d_bytes = ids.numel() * 8
        x = ops.philox_normal((1, nf, h, w, 4), seed, 0, 0, device=dev).float() * sch.init_noise_sigma
        bufs = {"x": x}
        for i, t in enumerate(sch.timesteps):
            xin = (bufs["x"] * float(np.float32(sch.input_scale(i)))).to(dt).permute(0, 4, 1, 2, 3).expand(2, -1, -1, -1, -1)
            e = self.unet(xin.contiguous(), torch.full((2,), t, device=dev), ctx).permute(0, 2, 3, 4, 1).float()
            bufs["eps"] = e[0:1] + float(np.float32(g)) * (e[1:2] - e[0:1])
            if sch.stochastic:
                bufs["noise"] = ops.philox_normal((1, nf, h, w, 4), seed, 1 + i, 0, device=dev).float()
            apply_plan_torch(sch.plan(i), bufs)
        lat = bufs["x"][0].to(dt).permute(0, 3, 1, 2)                           # [F, 4, h, w]
        img = torch.cat([self.vae(lat[i:i + 4]) for i in range(0, nf, 4)], 0).permute(0, 2, 3, 1)
        u8 = ((img.float() / 2 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8)
        out = u8.cpu().numpy()
        self.d2h_bytes = out.size
        return out

    def _run_native(self, st: dict, g: float) -> None:
        """``nf`` below is the number of frames THIS rank holds (all of them unless frame_parallel)."""
        import sys
        gemm_mod = sys.modules["arbius_b200.ops.gemm"]
        sch, nf = st["sch"], st["nfl"]
        P = self.fp.world if self.fp is not None else 1
        ctx = self.clip.forward(st["ids"])
        kvs = self.unet.precompute_context(ctx)
        temb = self.unet.time_projections(st["tvec"], nf)                      # [2*steps*F, sum(Cout)]
        f = st["f32"]
        ops.lincomb(f["x"], [(sch.init_noise_sigma, st["noise0"])])
        for i in range(len(sch.timesteps)):
            for b in range(2):
                ops.lincomb(None, [(1.0, f["x"])], dst_m=st["xin"][b], m_scale=sch.input_scale(i))
            if self.cfgp is None:
                with gemm_mod.batch_scale(P):          # every UNet GEMM's rows scale with the frames: canonical = all
                    e = self.unet.forward(st["xin"], temb[2 * nf * i:2 * nf * (i + 1)], kvs)
                ops.cfg_step(e[1], e[0], None, g=g, eps_out=f["eps"])
            else:
                self._cfg_parallel_step(st, i, temb, kvs, g)
            for dst, terms in sch.plan(i):
                ops.lincomb(f[dst], [(c, st["noise"][i] if s == "noise" else f[s]) for c, s in terms])
        ops.lincomb(None, [(1.0, f["x"])], dst_m=st["lat"], m_scale=1.0)
        chunks = []
        chunk = st["vae_chunk"]
        vscale = 1
        if P > 1 and chunk % P == 0:
            chunk, vscale = chunk // P, P              # the single-GPU chunk, cut over the ranks
        with gemm_mod.batch_scale(vscale):
            for i in range(0, nf, chunk):
                chunks.append(ops.finalize_u8(self.vae.forward(st["lat"][0, i:i + chunk])))
        st["u8"] = torch.cat(chunks, 0) if len(chunks) > 1 else chunks[0]
        if self.fp is not None:
            # every rank stores its frames into rank 0's clip buffer, then all meet: rank 0 owns the whole clip
            from ...parallel.frame_shard import _lib as fs_lib
            u8 = st["u8"].contiguous()
            nbytes = u8.numel()
            dst = self.fp.peer_ptr(0, st["clip_off"]) + self.fp.rank * nbytes
            ops.native.check(fs_lib().ab_peer_copy2d(dst, nbytes, u8.data_ptr(), nbytes, nbytes, 1,
                                                     ops.native.stream_ptr()), "ab_peer_copy2d")
            self.fp.barrier_replayable()
            st["u8"] = st["clip"]

    def _cfg_parallel_step(self, st: dict, i: int, temb, kvs, g: float) -> None:
        """This rank's guidance branch of step i, exchanged through the peer heap (see SD15Pipeline._cfg_parallel_step)."""
        import sys
        heap, r, par, nf = self.cfgp, self.cfgp.rank, i & 1, st["nfl"]
        gemm_mod = sys.modules["arbius_b200.ops.gemm"]
        with gemm_mod.batch_scale(2 * (self.fp.world if self.fp is not None else 1)):
            rows = temb[2 * nf * i + r * nf:2 * nf * i + (r + 1) * nf]
            e = self.unet.forward(st["xin"][r:r + 1], rows, [kv[r:r + 1] for kv in kvs])
        slot = st["ex"][par]
        slot.copy_(e)
        heap.barrier_replayable()
        own = slot.data_ptr()
        peer = heap.peer_ptr(1 - r, st["ex_off"] + par * slot.numel() * slot.element_size())
        ptr_c, ptr_u = (own, peer) if r == 1 else (peer, own)
        lib = ops.native.load()
        ops.native.check(lib.ab_cfg_step(ptr_c, ptr_u, None, None, None, None, None, st["f32"]["eps"].data_ptr(),
                                         slot.numel(), g, 0.0, 1.0, 0.0, 0.0, ops.native.dtype_code(slot.dtype),
                                         ops.native.stream_ptr()), "ab_cfg_step")

    def _native(self, ids, nf, steps, width, height, g, sched, seed) -> np.ndarray:
        dev, dt = self.device, self.dtype
        h, w = height // self.vae_scale, width // self.vae_scale
        key = (nf, steps, width, height, g, sched)
        P = self.fp.world if self.fp is not None else 1
        rank = self.fp.rank if self.fp is not None else 0
        if P > 1 and not frame_split_ok(nf, h, w, P, len(self.cfg.unet.block_out_channels)):
            # raised on every rank before the first barrier; the miner's backend then runs the task on one GPU
            raise UnsupportedSplit(f"{nf} frames of {h}x{w} latents do not shard over {P} ranks")
        nfl = nf // P
        st = self._graphs.get(key)
        if st is None:
            sch = make_scheduler(sched)
            sch.set_timesteps(steps)
            names = {"x", "eps", "tmp"} | set(sch.history)
            shape = (1, nfl, h, w, 4)
            st = dict(sch=sch, nf=nf, nfl=nfl, ids=torch.zeros((2, self.tokenizer.max_len), dtype=torch.int64, device=dev),
                      ids_host=torch.zeros((2, self.tokenizer.max_len), dtype=torch.int64).pin_memory(),
                      tvec=torch.tensor([t for t in sch.timesteps for _ in range(2)], dtype=torch.float32, device=dev),
                      noise0=torch.zeros(shape, dtype=torch.float32, device=dev),
                      noise=[torch.zeros(shape, dtype=torch.float32, device=dev) for _ in sch.timesteps] if sch.stochastic else [],
                      f32={n: torch.zeros(shape, dtype=torch.float32, device=dev) for n in names},
                      xin=torch.zeros((2, nfl, h, w, 4), dtype=dt, device=dev), lat=torch.zeros(shape, dtype=dt, device=dev),
                      out_host=torch.zeros((nf, height, width, 3), dtype=torch.uint8).pin_memory(),
                      vae_chunk=max(1, min(nf, (1 << 22) // max(1, height * width // 4))), graph=None, launches=0)
            if self.cfgp is not None:
                st["ex"], st["ex_off"] = self.cfgp.alloc((2,) + shape, dt)
            if self.fp is not None:
                from ...parallel.frame_shard import FrameShard
                ucfg = self.cfg.unet
                max_c = max(ucfg.block_out_channels[0], self.unet.transformer_in.inner)
                self.fp.reset()
                self._shard = FrameShard(self.fp, 2 if self.cfgp is None else 1, nfl, h * w, max_c, dt, ucfg.norm_groups)
                self.unet.set_frame_shard(self._shard)
                st["clip"], st["clip_off"] = self.fp.alloc((nf, height, width, 3), torch.uint8)
                self._graphs.clear()                   # the heap was laid out anew: older states' buffers are gone
            heaps = [hp for hp in (self.cfgp, self.fp) if hp is not None]
            if len(heaps) == 2:
                heaps[1].dry = getattr(heaps[1], "_local", False)      # one dry pass with BOTH groups' barriers off
            if heaps:
                try:
                    heaps[0].dry_run(lambda: self._run_native(st, g))
                finally:
                    heaps[-1].dry = False
            if self.use_graph:
                s = torch.cuda.Stream()
                s.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s):
                    self._run_native(st, g)
                torch.cuda.current_stream().wait_stream(s)
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                n0 = ops.native.launch_count()
                with ops.native.graph_capture(graph):
                    self._run_native(st, g)
                st["launches"] = ops.native.launch_count() - n0
                st["graph"] = graph
            self._graphs[key] = st
        st["ids_host"].copy_(ids)
        st["ids"].copy_(st["ids_host"], non_blocking=True)
        self.h2d_bytes = ids.numel() * 8
        lib = ops.native.load()
        n = nfl * h * w * 4
        off = rank * n                                  # global element index of this rank's first frame
        ops.native.check(lib.ab_philox_normal(st["noise0"].data_ptr(), n, seed, 0, off, ops.native.stream_ptr()), "philox")
        for i, nb in enumerate(st["noise"]):
            ops.native.check(lib.ab_philox_normal(nb.data_ptr(), n, seed, 1 + i, off, ops.native.stream_ptr()), "philox")
        if st["graph"] is not None:
            st["graph"].replay()
            self.last_launches = st["launches"] + 1 + len(st["noise"])
        else:
            n0 = ops.native.launch_count()
            self._run_native(st, g)
            self.last_launches = ops.native.launch_count() - n0 + 1 + len(st["noise"])
        if rank != 0:
            torch.cuda.current_stream().synchronize()
            self.d2h_bytes = 0
            return None
        st["out_host"].copy_(st["u8"], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self.d2h_bytes = st["out_host"].numel()
        return st["out_host"].numpy().copy()

    # ------------------------------------------------------------------------------------------
    def predict(self, inp: dict) -> bytes:
        """Hydrated template input (+ ``seed``) -> bytes of ``out-1.mp4``."""
        from ...codecs.mp4 import encode_mp4
        frames = self.generate(prompt=inp["prompt"], negative_prompt=inp.get("negative_prompt", "") or "",
                               num_frames=inp.get("num_frames"), num_inference_steps=int(inp.get("num_inference_steps", 50)),
                               width=inp.get("width"), height=inp.get("height"), guidance_scale=inp.get("guidance_scale"),
                               seed=int(inp.get("seed", 0)))
        if frames is None:                              # a frame-parallel rank other than 0: rank 0 has the clip
            return None
        return encode_mp4(frames, int(inp.get("fps", self.cfg.default_fps)))

    def predict_many(self, inputs) -> list:
        """A queue of hydrated inputs -> MP4 bytes in order; the MJPEG/MP4 encode of clip i (host, GIL released in the
        native JPEG encoder) runs on a worker thread while the GPU denoises clip i+1."""
        from concurrent.futures import ThreadPoolExecutor

        from ...codecs.mp4 import encode_mp4
        with ThreadPoolExecutor(max_workers=2) as pool:
            futs = []
            for inp in inputs:
                frames = self.generate(prompt=inp["prompt"], negative_prompt=inp.get("negative_prompt", "") or "",
                                       num_frames=inp.get("num_frames"),
                                       num_inference_steps=int(inp.get("num_inference_steps", 50)),
                                       width=inp.get("width"), height=inp.get("height"),
                                       guidance_scale=inp.get("guidance_scale"), seed=int(inp.get("seed", 0)))
                futs.append(None if frames is None else
                            pool.submit(encode_mp4, frames, int(inp.get("fps", self.cfg.default_fps))))
            return [None if f is None else f.result() for f in futs]
