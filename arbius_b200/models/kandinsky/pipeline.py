"""Kandinsky 2.1 text-to-image pipeline: prompt -> PNG bytes (the model the reference miner ships enabled).

Mirrors ``get_kandinsky2('cuda', task_type='text2img', model_version='2.1').generate_text2img`` of the container
named in templates/kandinsky2.json (reference: miner/src/index.ts:781-877 Kandinsky2Model; container defaults in
docs/src/pages/register-model.mdx:124-219: 100 steps, guidance 4, ``p_sampler``, prior_cf_scale 4, prior_steps "5"):

  1. XLM-R text encoder on [negative "", prompt] -> token features + pooled embedding
  2. diffusion prior: 5 DDIM steps (cosine schedule, x0 parameterisation, CFG 4) -> CLIP image embedding
  3. UNet: 100 ancestral steps at batch 2 (uncond uses a zero image embedding), guidance 4, 768^2 -> 96^2 latent
  4. MoVQ decode -> uint8 -> deterministic PNG
The template only exposes prompt / width / height; everything else is the container's fixed default.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from ... import ops
from ..schedulers import DDIM, apply_plan_torch, make_scheduler
from ..sd15.unet_ref import init_weights_
from ..tokenizer import HashTokenizer
from .ref import (DiffusionPrior, KandinskyUNet, KUNetConfig, MoVQConfig, MoVQDecoder, PriorConfig, TextEncoderXLMR,
                  XlmrConfig)


@dataclass
class Kandinsky2Config:
    unet: KUNetConfig = field(default_factory=KUNetConfig)
    prior: PriorConfig = field(default_factory=PriorConfig)
    text: XlmrConfig = field(default_factory=XlmrConfig)
    movq: MoVQConfig = field(default_factory=MoVQConfig)
    steps: int = 100
    guidance: float = 4.0
    sampler: str = "p_sampler"
    prior_steps: int = 5
    prior_guidance: float = 4.0

    @staticmethod
    def tiny() -> "Kandinsky2Config":
        return Kandinsky2Config(KUNetConfig.tiny(), PriorConfig.tiny(), XlmrConfig.tiny(), MoVQConfig.tiny(), steps=3,
                                prior_steps=2)


class Kandinsky2Pipeline:
    def __init__(self, cfg: Kandinsky2Config = Kandinsky2Config(), backend: str = "native", device: str = "cuda",
                 dtype=torch.float16, seed: int = 0, model_dir: Optional[str] = None, use_graph: bool = True,
                 cfg_parallel=None, patch_parallel=None):
        """``cfg_parallel``: a 2-rank ``PeerHeap`` — the two classifier-free-guidance branches of every decoder-UNet step
        run on two GPUs and exchange their predictions through peer memory (as in ``SD15Pipeline``); text encoder,
        prior and MoVQ are small and run whole on both. Same bytes as one GPU.
        ``patch_parallel``: a 2/4/8-rank ``PeerHeap`` — ONE image over those GPUs, the decoder UNet split into row bands
        (parallel/patch.py: GroupNorm moments, conv halo rows and attention K/V cross ranks through peer memory with
        in-kernel flag waits). Text encoder, prior and MoVQ run whole on every rank. Same bytes as one GPU."""
        self.cfg, self.backend, self.device, self.dtype = cfg, backend, torch.device(device), dtype
        self.cfgp = cfg_parallel
        assert cfg_parallel is None or (cfg_parallel.world == 2 and backend == "native"), "cfg_parallel needs a 2-rank heap"
        self.pp = None
        if patch_parallel is not None:
            assert backend == "native" and cfg_parallel is None
            from ...parallel.patch import PatchContext
            self.pp = PatchContext(patch_parallel, dtype)
        from ..tokenizer import load_xlmr_tokenizer
        self.tokenizer = load_xlmr_tokenizer(model_dir, cfg.text.vocab_size, 77)
        with torch.device(device):
            unet, prior = KandinskyUNet(cfg.unet), DiffusionPrior(cfg.prior, text_dim=cfg.text.hidden)
            text, movq = TextEncoderXLMR(cfg.text), MoVQDecoder(cfg.movq)
        mods = (unet, prior, text, movq)
        if model_dir:
            from pathlib import Path
            ck = Path(model_dir)
            from . import keymap
            for mod, name in zip(mods, ("decoder_fp16.ckpt", "prior_fp16.ckpt", "text_encoder.ckpt", "movq_final.ckpt")):
                sd = torch.load(ck / name, map_location="cpu", weights_only=True)
                sd = sd.get("state_dict", sd)
                # the files of kasumi-1/Kandinsky-2 carry upstream names and per-head interleaved attention projections
                # (keymap.py); files saved from these modules load as they are
                if name.startswith("decoder") and not sd.get("__arbius_layout__"):
                    sd = keymap.map_unet(sd, cfg.unet.head_dim)
                elif name.startswith("prior") and keymap.is_upstream_prior(sd):
                    sd = keymap.map_prior(sd, cfg.prior.heads)
                elif name.startswith("text_encoder") and keymap.is_upstream_text_encoder(sd):
                    sd = keymap.map_text_encoder(sd, cfg.text.max_positions)
                elif name.startswith("movq") and keymap.is_upstream_movq(sd):
                    sd = keymap.map_movq(sd, len(cfg.movq.ch_mult))
                sd.pop("__arbius_layout__", None)
                mod.load_state_dict(sd, strict=True)
        else:
            for i, m in enumerate(mods):
                init_weights_(m, seed + 20 + i)
        for m in mods:
            m.to(dtype).eval().requires_grad_(False)
        self.vae_scale = 2 ** (len(cfg.movq.ch_mult) - 1)
        self.use_graph = use_graph and backend == "native"
        self._graphs: Dict[Tuple, dict] = {}
        if backend == "eager":
            self.unet, self.prior, self.text, self.movq = mods
        else:
            from .native import NativeKandinskyUNet, NativeMoVQDecoder, NativePrior, NativeXlmr
            usd = dict(unet.state_dict())
            usd["out.2.weight"], usd["out.2.bias"] = usd["out.2.weight"][:4], usd["out.2.bias"][:4]   # eps half only
            self.unet = NativeKandinskyUNet(usd, cfg.unet, dtype, device)
            self.prior = NativePrior(prior.state_dict(), cfg.prior, dtype, device)
            self.text = NativeXlmr(text.state_dict(), cfg.text, dtype, device)
            self.movq = NativeMoVQDecoder(movq.state_dict(), cfg.movq, dtype, device)
            del mods, unet, prior, text, movq
        self.h2d_bytes = self.d2h_bytes = 0
        self.last_launches = 0

    # ---------------------------------------------------------------------------------------- shared pieces
    def _prior_sched(self) -> DDIM:
        s = DDIM(beta_schedule="cosine", steps_offset=0)
        s.set_timesteps(self.cfg.prior_steps)
        return s

    def _ids(self, prompt: str) -> torch.Tensor:
        return torch.tensor([self.tokenizer.encode(""), self.tokenizer.encode(prompt)], dtype=torch.int64)

    @torch.no_grad()
    def generate(self, prompt: str, width: int = 768, height: int = 768, seed: int = 0) -> np.ndarray:
        ids = self._ids(prompt)
        if self.backend == "eager":
            return self._eager(ids, width, height, seed)
        return self._native(ids, width, height, seed)

    # ---------------------------------------------------------------------------------------- eager oracle
    def _eager(self, ids, width, height, seed) -> np.ndarray:
        c, dev, dt = self.cfg, self.device, self.dtype
        h, w = height // self.vae_scale, width // self.vae_scale
        tokens, pooled = self.text(ids.to(dev))
        self.h2d_bytes = ids.numel() * 8
        # ---- prior ----
        ps = self._prior_sched()
        bufs = {"x": ops.philox_normal((1, c.prior.clip_dim), seed, 100, 0, device=dev).float()}
        for i, t in enumerate(ps.timesteps):
            a_t = float(ps.alphas_cumprod[int(t)])
            xin = bufs["x"].to(dt).expand(2, -1).contiguous()
            x0 = self.prior(xin, torch.full((2,), t, device=dev), tokens, pooled).float()
            x0 = x0[0:1] + float(np.float32(c.prior_guidance)) * (x0[1:2] - x0[0:1])
            bufs["eps"] = (bufs["x"] - float(np.float32(math.sqrt(a_t))) * x0) * float(np.float32(1.0 / math.sqrt(1 - a_t)))
            apply_plan_torch(ps.plan(i), bufs)
        image_emb = torch.cat([torch.zeros_like(bufs["x"]), bufs["x"]], 0).to(dt)
        # ---- decoder UNet ----
        sch = make_scheduler(c.sampler, beta_schedule="linear", beta_start=0.00085, beta_end=0.012)
        sch.set_timesteps(c.steps)
        bufs = {"x": ops.philox_normal((1, h, w, 4), seed, 0, 0, device=dev).float() * sch.init_noise_sigma}
        for i, t in enumerate(sch.timesteps):
            xin = bufs["x"].to(dt).permute(0, 3, 1, 2).expand(2, -1, -1, -1).contiguous()
            out = self.unet(xin, torch.full((2,), t, device=dev), tokens, pooled, image_emb)
            e = out[:, :4].permute(0, 2, 3, 1).float()
            bufs["eps"] = e[0:1] + float(np.float32(c.guidance)) * (e[1:2] - e[0:1])
            if sch.stochastic:
                bufs["noise"] = ops.philox_normal((1, h, w, 4), seed, 1 + i, 0, device=dev).float()
            apply_plan_torch(sch.plan(i), bufs)
        img = self.movq(bufs["x"].to(dt).permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
        u8 = ((img.float() / 2 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8)[0]
        out = u8.cpu().numpy()
        self.d2h_bytes = out.size
        return out

    # ---------------------------------------------------------------------------------------- native
    def _run_native(self, st: dict) -> None:
        c, dt = self.cfg, self.dtype
        f = st["f32"]
        tokens, pooled = self.text.forward(st["ids"])
        # ---- prior (x0 parameterisation -> eps -> DDIM plan), batch 2 for CFG ----
        ps = st["ps"]
        tt, te = self.prior.text_part(tokens, pooled)
        ops.lincomb(f["px"], [(1.0, st["pnoise"])])
        for i, t in enumerate(ps.timesteps):
            a_t = float(ps.alphas_cumprod[int(t)])
            for b in range(2):
                ops.lincomb(None, [(1.0, f["px"])], dst_m=st["pxin"][b], m_scale=1.0)
            x0 = self.prior.forward(st["pxin"], st["ptvec"][2 * i:2 * i + 2], tt, te)
            ops.cfg_step(x0[1], x0[0], None, g=c.prior_guidance, eps_out=f["px0"])
            ops.lincomb(f["peps"], [(1.0 / math.sqrt(1 - a_t), f["px"]), (-math.sqrt(a_t) / math.sqrt(1 - a_t), f["px0"])])
            for dst, terms in ps.plan(i):
                ops.lincomb(f["p" + dst], [(co, f["p" + s]) for co, s in terms])
        ops.lincomb(None, [(1.0, f["px"])], dst_m=st["image_emb"][1], m_scale=1.0)       # row 0 stays zero (uncond)
        # ---- decoder ----
        sch = st["sch"]
        cond = self.unet.prepare(tokens, pooled, st["image_emb"], st["h"] * st["w"])
        ops.lincomb(f["x"], [(sch.init_noise_sigma, st["noise0"])])
        for i in range(len(sch.timesteps)):
            for b in range(2):
                ops.lincomb(None, [(1.0, f["x"])], dst_m=st["xin"][b], m_scale=sch.input_scale(i))
            emb_all = self.unet.embeddings(st["tvec"][2 * i:2 * i + 2], cond)
            if self.pp is not None:
                e = self.unet.forward_patch(self.pp, st["xin"], emb_all)
                ops.cfg_step(e[1], e[0], None, g=c.guidance, eps_out=f["eps"])
            elif self.cfgp is None:
                e = self.unet.forward(st["xin"], emb_all)
                ops.cfg_step(e[1], e[0], None, g=c.guidance, eps_out=f["eps"])
            else:
                self._cfg_parallel_step(st, i, emb_all)
            for dst, terms in sch.plan(i):
                ops.lincomb(f[dst], [(co, st["noise"][i] if s == "noise" else f[s]) for co, s in terms])
        ops.lincomb(None, [(1.0, f["x"])], dst_m=st["lat"], m_scale=1.0)
        st["u8"] = ops.finalize_u8(self.movq.forward(st["lat"]))

    def _cfg_parallel_step(self, st: dict, i: int, emb_all) -> None:
        """This rank's guidance branch of decoder step i; predictions cross through two parity slots in the peer heap
        behind one device barrier per step (see SD15Pipeline._cfg_parallel_step)."""
        import sys
        heap, r, par = self.cfgp, self.cfgp.rank, i & 1
        gemm_mod = sys.modules["arbius_b200.ops.gemm"]
        with gemm_mod.batch_scale(2):                 # split-K choices of the undivided batch of two
            e = self.unet.forward(st["xin"][r:r + 1], emb_all[r:r + 1], sample_offset=r)
        slot = st["ex"][par]
        slot.copy_(e)
        heap.barrier_replayable()
        own = slot.data_ptr()
        peer = heap.peer_ptr(1 - r, st["ex_off"] + par * slot.numel() * slot.element_size())
        ptr_c, ptr_u = (own, peer) if r == 1 else (peer, own)
        lib = ops.native.load()
        ops.native.check(lib.ab_cfg_step(ptr_c, ptr_u, None, None, None, None, None, st["f32"]["eps"].data_ptr(),
                                         slot.numel(), self.cfg.guidance, 0.0, 1.0, 0.0, 0.0,
                                         ops.native.dtype_code(slot.dtype), ops.native.stream_ptr()), "ab_cfg_step")

    def _native(self, ids, width, height, seed) -> np.ndarray:
        c, dev, dt = self.cfg, self.device, self.dtype
        h, w = height // self.vae_scale, width // self.vae_scale
        key = (width, height)
        st = self._graphs.get(key)
        if st is None:
            sch = make_scheduler(c.sampler, beta_schedule="linear", beta_start=0.00085, beta_end=0.012)
            sch.set_timesteps(c.steps)
            ps = self._prior_sched()
            shape, pshape = (1, h, w, 4), (1, c.prior.clip_dim)
            names = {"x", "eps", "tmp"} | set(sch.history)
            f32 = {n: torch.zeros(shape, dtype=torch.float32, device=dev) for n in names}
            for n in ("px", "peps", "px0", "ptmp"):
                f32[n] = torch.zeros(pshape, dtype=torch.float32, device=dev)
            st = dict(sch=sch, ps=ps, h=h, w=w, f32=f32,
                      ids=torch.zeros((2, 77), dtype=torch.int64, device=dev),
                      ids_host=torch.zeros((2, 77), dtype=torch.int64).pin_memory(),
                      tvec=torch.tensor([t for t in sch.timesteps for _ in range(2)], dtype=torch.float32, device=dev),
                      ptvec=torch.tensor([t for t in ps.timesteps for _ in range(2)], dtype=torch.float32, device=dev),
                      noise0=torch.zeros(shape, dtype=torch.float32, device=dev),
                      noise=[torch.zeros(shape, dtype=torch.float32, device=dev) for _ in sch.timesteps] if sch.stochastic else [],
                      pnoise=torch.zeros(pshape, dtype=torch.float32, device=dev),
                      pxin=torch.zeros((2, c.prior.clip_dim), dtype=dt, device=dev),
                      image_emb=torch.zeros((2, c.prior.clip_dim), dtype=dt, device=dev),
                      xin=torch.zeros((2, h, w, 4), dtype=dt, device=dev), lat=torch.zeros(shape, dtype=dt, device=dev),
                      out_host=torch.zeros((height, width, 3), dtype=torch.uint8).pin_memory(), graph=None, launches=0)
            if self.cfgp is not None:
                st["ex"], st["ex_off"] = self.cfgp.alloc((2, 1, h, w, 4), dt)      # same offset on both ranks
            if self.cfgp is not None:
                self.cfgp.dry_run(lambda: self._run_native(st))
            if self.pp is not None:
                self.pp.heap.dry_run(lambda: self._run_native(st))
            if self.use_graph:
                s = torch.cuda.Stream()
                s.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s):
                    self._run_native(st)
                torch.cuda.current_stream().wait_stream(s)
                torch.cuda.synchronize()
                graph = torch.cuda.CUDAGraph()
                n0 = ops.native.launch_count()
                with ops.native.graph_capture(graph):
                    self._run_native(st)
                st["launches"] = ops.native.launch_count() - n0
                st["graph"] = graph
            self._graphs[key] = st
        st["ids_host"].copy_(ids)
        st["ids"].copy_(st["ids_host"], non_blocking=True)
        self.h2d_bytes = ids.numel() * 8
        lib, sp = ops.native.load(), ops.native.stream_ptr
        ops.native.check(lib.ab_philox_normal(st["noise0"].data_ptr(), h * w * 4, seed, 0, 0, sp()), "philox")
        ops.native.check(lib.ab_philox_normal(st["pnoise"].data_ptr(), c.prior.clip_dim, seed, 100, 0, sp()), "philox")
        for i, nb in enumerate(st["noise"]):
            ops.native.check(lib.ab_philox_normal(nb.data_ptr(), h * w * 4, seed, 1 + i, 0, sp()), "philox")
        if st["graph"] is not None:
            st["graph"].replay()
            self.last_launches = st["launches"] + 2 + len(st["noise"])
        else:
            n0 = ops.native.launch_count()
            self._run_native(st)
            self.last_launches = ops.native.launch_count() - n0 + 2 + len(st["noise"])
        st["out_host"].copy_(st["u8"][0], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self.d2h_bytes = st["out_host"].numel()
        return st["out_host"].numpy().copy()

    def predict(self, inp: dict) -> bytes:
        from ...codecs.png import encode_png
        return encode_png(self.generate(inp["prompt"], int(inp.get("width", 768)), int(inp.get("height", 768)),
                                        int(inp.get("seed", 0))))
