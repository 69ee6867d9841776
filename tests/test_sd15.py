"""SD-1.5 models: scheduler plans, eager oracle on CPU, native engines vs the oracle on GPU."""
import math

import numpy as np
import pytest
import torch

from arbius_b200.models.schedulers import SCHEDULERS, apply_plan_torch, make_scheduler
from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline, build_modules
from arbius_b200.models.tokenizer import HashTokenizer

TEMPLATE_SCHEDULERS = ["DDIM", "K_EULER", "DPMSolverMultistep", "K_EULER_ANCESTRAL", "PNDM", "KLMS"]


@pytest.mark.parametrize("name", TEMPLATE_SCHEDULERS + ["p_sampler"])
@pytest.mark.parametrize("steps", [20, 50])
def test_scheduler_recovers_point_mass(name, steps):
    """For data concentrated at x0 the exact eps is (x - sqrt(ab) x0)/sqrt(1-ab) (VP) or (x - x0)/sigma
    (k-diffusion); every sampler must walk the noise back to x0."""
    sch = make_scheduler(name)
    sch.set_timesteps(steps)
    x0 = torch.tensor([0.7, -1.3, 0.25, 2.0])
    g = torch.Generator().manual_seed(0)
    bufs = {"x": torch.randn(4, generator=g) * sch.init_noise_sigma}
    kdiff = hasattr(sch, "sigmas") and name != "DPMSolverMultistep"
    for i, t in enumerate(sch.timesteps):
        if kdiff:
            sigma = sch.sigmas[i]
            bufs["eps"] = (bufs["x"] - x0) / sigma
        else:
            ab = float(np.interp(t, np.arange(sch.T), sch.alphas_cumprod))
            bufs["eps"] = (bufs["x"] - math.sqrt(ab) * x0) / math.sqrt(1 - ab)
        bufs["noise"] = torch.randn(4, generator=g)
        apply_plan_torch(sch.plan(i), bufs)
    # DDIM/PNDM stop at alphas_cumprod[0] (set_alpha_to_one=False): ~0.03 sigma of noise legitimately remains
    tol = 0.2 if sch.stochastic else 0.1
    assert torch.allclose(bufs["x"], x0, atol=tol), (name, bufs["x"])


def test_scheduler_registry_and_shapes():
    for n in TEMPLATE_SCHEDULERS:
        assert n in SCHEDULERS
    s = make_scheduler("PNDM")
    s.set_timesteps(20)
    assert len(s.timesteps) == 21 and s.timesteps[1] == s.timesteps[2]
    d = make_scheduler("DPMSolverMultistep")
    d.set_timesteps(20)
    assert len(d.timesteps) == 20 and d.timesteps[0] == 999 and d.sigmas[-1] == 0
    for dst, terms in d.plan(19):
        assert dst == "x" and len(terms) == 2
    with pytest.raises(ValueError):
        make_scheduler("nope")


def test_hash_tokenizer_framing():
    t = HashTokenizer()
    ids = t.encode("Arbius test cat, 4k!")
    assert len(ids) == 77 and ids[0] == 49406 and ids[-1] == 49407
    assert ids == t.encode("arbius  test cat, 4k!") and ids != t.encode("arbius test dog, 4k!")
    assert t.encode("")[:2] == [49406, 49407]


def test_eager_pipeline_cpu_is_deterministic():
    pipe = SD15Pipeline(SD15Config.tiny(), backend="eager", device="cpu", dtype=torch.float32)
    a = pipe.generate("arbius test cat", "", 64, 64, 3, 7.5, "DPMSolverMultistep", seed=1337)
    b = pipe.generate("arbius test cat", "", 64, 64, 3, 7.5, "DPMSolverMultistep", seed=1337)
    c = pipe.generate("arbius test cat", "", 64, 64, 3, 7.5, "DPMSolverMultistep", seed=1338)
    assert a.shape == (64, 64, 3) and a.dtype == np.uint8
    assert (a == b).all() and (a != c).any()
    png = pipe.predict({"prompt": "x", "negative_prompt": "", "width": 64, "height": 64, "num_inference_steps": 2,
                        "guidance_scale": 7, "scheduler": "K_EULER_ANCESTRAL", "seed": 5})
    assert png[:8] == b"\x89PNG\r\n\x1a\n"
    # the queue API returns the same bytes, in order
    inputs = [{"prompt": "x", "negative_prompt": "", "width": 64, "height": 64, "num_inference_steps": 2,
               "guidance_scale": 7, "scheduler": "K_EULER_ANCESTRAL", "seed": s} for s in (5, 6, 5)]
    many = pipe.predict_many(inputs)
    assert many[0] == png and many[2] == png and many[1] != png


# ------------------------------------------------------------------ GPU
def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16])
def test_native_engines_match_eager_fp32(dt):
    from arbius_b200.models.sd15.native import NativeClipText, NativeUNet, NativeVaeDecoder
    cfg = SD15Config.tiny()
    unet, clip, vae = build_modules(cfg, "cuda", torch.float32, seed=3)
    tol = 0.03 if dt == torch.float16 else 0.12
    # CLIP
    ids = torch.randint(0, cfg.clip.vocab_size, (2, 77), device="cuda")
    ref = clip(ids)
    got = NativeClipText(clip.state_dict(), cfg.clip, dt, "cuda").forward(ids).float()
    assert (got - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())
    # UNet
    nu = NativeUNet(unet.state_dict(), cfg.unet, dt, "cuda")
    x = torch.randn(2, 4, 16, 16, device="cuda")
    ctx = torch.randn(2, 77, cfg.unet.cross_attention_dim, device="cuda")
    t = torch.tensor([500.0, 500.0], device="cuda")
    ref = _nhwc(unet(x, t, ctx))
    kvs = nu.precompute_context(ctx.to(dt))
    got = nu.forward(_nhwc(x).to(dt), nu.time_projections(t), kvs).float()
    assert (got - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())
    # VAE
    z = torch.randn(1, 4, 8, 8, device="cuda") * 0.18215
    ref = _nhwc(vae(z))
    got = NativeVaeDecoder(vae.state_dict(), cfg.vae, dt, "cuda").forward(_nhwc(z).to(dt)).float()
    assert (got - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item())


@pytest.mark.gpu
@pytest.mark.parametrize("sched", ["DPMSolverMultistep", "K_EULER_ANCESTRAL", "PNDM"])
def test_native_pipeline_matches_eager_and_is_bit_reproducible(sched):
    cfg = SD15Config.tiny()
    kw = dict(prompt="arbius test cat", negative_prompt="blurry", width=64, height=64, num_inference_steps=3,
              guidance_scale=7.5, scheduler=sched, seed=1337)
    eager = SD15Pipeline(cfg, backend="eager", device="cuda", dtype=torch.float32).generate(**kw)
    nat = SD15Pipeline(cfg, backend="native", device="cuda", dtype=torch.float16, use_graph=True)
    a = nat.generate(**kw)
    b = nat.generate(**kw)                       # graph replay
    assert (a == b).all(), "native output must be bit-identical run to run"
    diff = np.abs(a.astype(np.int32) - eager.astype(np.int32))
    assert diff.mean() < 2.0 and diff.max() <= 40, (diff.mean(), diff.max())
    nograph = SD15Pipeline(cfg, backend="native", device="cuda", dtype=torch.float16, use_graph=False).generate(**kw)
    assert (a == nograph).all(), "graph replay and eager launch must agree bit for bit"
    kw2 = dict(kw, seed=4242)
    assert (nat.generate(**kw2) != a).any()
    assert nat.last_launches > 0


@pytest.mark.gpu
@pytest.mark.parametrize("sched", ["DPMSolverMultistep", "K_EULER_ANCESTRAL"])
def test_task_batch_gives_each_task_the_bytes_it_gets_alone(sched):
    """Independent tasks that share one graph replay (UNet batch 2*n) must come out byte for byte as when they run alone:
    split-K choices are pinned to the lone task's, everything else is per-sample."""
    pipe = SD15Pipeline(SD15Config.tiny(), backend="native", device="cuda", dtype=torch.float16, use_graph=True)
    inputs = [{"prompt": f"prompt number {s}", "negative_prompt": "blurry" if s % 2 else "", "width": 64, "height": 64,
               "num_inference_steps": 3, "guidance_scale": 7.5, "scheduler": sched, "seed": 100 + s} for s in range(5)]
    alone = [pipe.predict(i) for i in inputs]
    assert len(set(alone)) == 5
    for tb in (2, 4):
        assert pipe.predict_many(inputs, task_batch=tb) == alone, f"task_batch={tb}"
    # a task with different settings in the middle of the queue breaks the group, not the result
    odd = dict(inputs[1], num_inference_steps=2)
    mixed = pipe.predict_many([inputs[0], odd, inputs[2], inputs[3]], task_batch=2)
    assert mixed[0] == alone[0] and mixed[2] == alone[2] and mixed[3] == alone[3] and mixed[1] == pipe.predict(odd)


@pytest.mark.gpu
def test_full_size_unet_step_matches_eager_fp32():
    """One UNet step of the REAL architecture (860 M parameters, 64x64 latent, batch 2) on the native fp16 kernels against
    the eager fp32 module with the same random weights: relative error of the noise prediction below 2 % of its range
    (fp16 storage of 60 layers; the tiny-config tests cannot see shape-dependent faults: tile choices, split-K, tails)."""
    from arbius_b200.models.sd15.native import NativeUNet
    cfg = SD15Config()
    unet, _, _ = build_modules(cfg, "cuda", torch.float32, seed=3)
    nu = NativeUNet(unet.state_dict(), cfg.unet, torch.float16, "cuda")
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(2, 4, 64, 64, device="cuda", generator=g)
    ctx = torch.randn(2, 77, cfg.unet.cross_attention_dim, device="cuda", generator=g)
    t = torch.tensor([500.0, 500.0], device="cuda")
    with torch.no_grad():
        ref = _nhwc(unet(x, t, ctx))
    got = nu.forward(_nhwc(x).half(), nu.time_projections(t), nu.precompute_context(ctx.half())).float()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert torch.isfinite(got).all() and err <= 0.02 * max(scale, 1.0), (err, scale)
    again = nu.forward(_nhwc(x).half(), nu.time_projections(t), nu.precompute_context(ctx.half())).float()
    assert torch.equal(got, again)
