"""Video UNet kernels and engine: temporal conv / temporal attention vs torch, NativeUNet3D vs the eager oracle."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _close(got, ref, tol):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs().max().item()
    assert torch.isfinite(got).all()
    assert err <= tol * max(1.0, ref.abs().max().item()), f"max err {err} (ref max {ref.abs().max().item()})"


@pytest.mark.parametrize("b,f,hw,c,co", [(2, 24, 144, 64, 64), (1, 16, 1024, 320, 320), (2, 8, 60, 128, 64)])
def test_temporal_conv_matches_conv3d(b, f, hw, c, co):
    from arbius_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(b, f, hw, c, device="cuda", dtype=torch.float16)
    w = torch.randn(co, c, 3, 1, 1, device="cuda", dtype=torch.float16) / (3 * c) ** 0.5
    bias = torch.randn(co, device="cuda", dtype=torch.float16)
    res = torch.randn(b, f, hw, co, device="cuda", dtype=torch.float16)
    y = ops.conv_temporal(x, ops.pack_temporal_conv_weight(w), bias, residual=res)
    xr = x.float().permute(0, 3, 1, 2).unsqueeze(-1)                      # [B, C, F, HW, 1]
    ref = F.conv3d(xr, w.float(), bias.float(), padding=(1, 0, 0)).squeeze(-1).permute(0, 2, 3, 1) + res.float()
    _close(y, ref, 4e-3)


@pytest.mark.parametrize("b,f,hw,h,d", [(2, 24, 144, 5, 64), (1, 16, 256, 2, 32), (2, 24, 37, 8, 64), (1, 8, 50, 4, 64)])
def test_temporal_attention_matches_sdpa(b, f, hw, h, d):
    from arbius_b200 import ops
    torch.manual_seed(1)
    c = h * d
    qkv = torch.randn(b, f, hw, 3 * c, device="cuda", dtype=torch.float16)
    q, k, v = (qkv[..., i * c:(i + 1) * c] for i in range(3))
    out = ops.temporal_attn(q, k, v, h)

    def seq(t):   # [B, F, HW, C] -> [B*HW, h, F, d]
        return t.float().permute(0, 2, 1, 3).reshape(b * hw, f, h, d).transpose(1, 2)
    ref = F.scaled_dot_product_attention(seq(q), seq(k), seq(v)).transpose(1, 2).reshape(b, hw, f, c).permute(0, 2, 1, 3)
    _close(out, ref, 4e-3)
    assert torch.equal(out, ops.temporal_attn(q, k, v, h))


def test_native_unet3d_matches_eager():
    from arbius_b200.models.sd15.unet_ref import init_weights_
    from arbius_b200.models.video.native import NativeUNet3D
    from arbius_b200.models.video.unet3d_ref import UNet3DConditionModel, UNet3DConfig
    cfg = UNet3DConfig.tiny()
    with torch.device("cuda"):
        ref_m = UNet3DConditionModel(cfg)
    init_weights_(ref_m, 5).eval().requires_grad_(False)
    nat = NativeUNet3D(ref_m.state_dict(), cfg, torch.float16, "cuda")
    torch.manual_seed(2)
    b, f, hh, ww = 2, 8, 16, 16
    x = torch.randn(b, 4, f, hh, ww, device="cuda")
    ctx = torch.randn(b, 77, cfg.cross_attention_dim, device="cuda")
    t = torch.tensor([400.0, 400.0], device="cuda")
    with torch.no_grad():
        ref = ref_m(x, t, ctx).permute(0, 2, 3, 4, 1)                     # [B, F, H, W, 4]
    kvs = nat.precompute_context(ctx.half())
    got = nat.forward(x.permute(0, 2, 3, 4, 1).contiguous().half(), nat.time_projections(t, f), kvs)
    _close(got, ref, 0.04)


def test_native_video_pipeline_matches_eager_and_is_reproducible():
    import numpy as np
    from arbius_b200.models.video.pipeline import TextToVideoConfig, TextToVideoPipeline
    cfg = TextToVideoConfig.tiny()
    kw = dict(prompt="a cat walking", negative_prompt="blurry", num_frames=8, num_inference_steps=3, width=32, height=32,
              guidance_scale=9.0, scheduler="DPMSolverMultistep", seed=11)
    eager = TextToVideoPipeline(cfg, backend="eager", device="cuda", dtype=torch.float32).generate(**kw)
    nat = TextToVideoPipeline(cfg, backend="native", device="cuda", dtype=torch.float16)
    a, b = nat.generate(**kw), nat.generate(**kw)
    assert a.shape == (8, 32, 32, 3) and (a == b).all()
    diff = np.abs(a.astype(np.int32) - eager.astype(np.int32))
    assert diff.mean() < 2.5, (diff.mean(), diff.max())
    assert nat.last_launches > 0
