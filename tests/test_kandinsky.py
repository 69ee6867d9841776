"""Kandinsky 2.1: eager pipeline on CPU; native engines and pipeline vs the eager oracle on GPU."""
import numpy as np
import pytest
import torch

from arbius_b200.models.kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
from arbius_b200.models.kandinsky.ref import (DiffusionPrior, KandinskyUNet, KUNetConfig, MoVQConfig, MoVQDecoder,
                                              PriorConfig, TextEncoderXLMR, XlmrConfig)
from arbius_b200.models.sd15.unet_ref import init_weights_


def test_eager_pipeline_cpu_deterministic():
    pipe = Kandinsky2Pipeline(Kandinsky2Config.tiny(), backend="eager", device="cpu", dtype=torch.float32)
    a = pipe.generate("arbius test cat", 32, 32, seed=1337)
    b = pipe.generate("arbius test cat", 32, 32, seed=1337)
    assert a.shape == (32, 32, 3) and a.dtype == np.uint8 and (a == b).all()
    assert (pipe.generate("arbius test cat", 32, 32, seed=1) != a).any()
    png = pipe.predict({"prompt": "arbius test cat", "width": 32, "height": 32, "seed": 1337})
    assert png[:4] == b"\x89PNG"


def test_unet_structure_matches_guided_diffusion_layout():
    with torch.device("meta"):
        u = KandinskyUNet()
    keys = set(u.state_dict())
    assert "input_blocks.0.0.weight" in keys and "middle_block.1.encoder_kv.weight" in keys
    assert "output_blocks.3.2.in_layers.2.weight" in keys           # resblock-up at the end of the lowest level
    assert u.state_dict()["out.2.weight"].shape == (8, 384, 3, 3)
    assert abs(sum(p.numel() for p in u.parameters()) / 1e9 - 1.2) < 0.1


def _close(got, ref, tol):
    got, ref = got.float(), ref.float()
    assert torch.isfinite(got).all()
    err = (got - ref).abs().max().item()
    assert err <= tol * max(1.0, ref.abs().max().item()), err


def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


@pytest.mark.gpu
def test_native_engines_match_eager():
    from arbius_b200.models.kandinsky.native import NativeKandinskyUNet, NativeMoVQDecoder, NativePrior, NativeXlmr
    torch.manual_seed(0)
    dt = torch.float16
    with torch.device("cuda"):
        unet, prior = KandinskyUNet(KUNetConfig.tiny()), DiffusionPrior(PriorConfig.tiny(), text_dim=64)
        text, movq = TextEncoderXLMR(XlmrConfig.tiny()), MoVQDecoder(MoVQConfig.tiny())
    for i, m in enumerate((unet, prior, text, movq)):
        init_weights_(m, i).eval().requires_grad_(False)
    ids = torch.randint(0, 1000, (2, 77), device="cuda")
    tok_ref, pooled_ref = text(ids)
    tok, pooled = NativeXlmr(text.state_dict(), XlmrConfig.tiny(), dt, "cuda").forward(ids)
    _close(tok, tok_ref, 0.03)
    _close(pooled, pooled_ref, 0.03)
    x = torch.randn(2, 64, device="cuda")
    t = torch.tensor([300.0, 300.0], device="cuda")
    np_ = NativePrior(prior.state_dict(), PriorConfig.tiny(), dt, "cuda")
    tt, te = np_.text_part(tok_ref.half(), pooled_ref.half())
    _close(np_.forward(x.half(), t, tt, te), prior(x, t, tok_ref, pooled_ref), 0.03)
    img = torch.randn(2, 64, device="cuda")
    lat = torch.randn(2, 4, 16, 16, device="cuda")
    ref = _nhwc(unet(lat, t, tok_ref, pooled_ref, img))
    nu = NativeKandinskyUNet(unet.state_dict(), KUNetConfig.tiny(), dt, "cuda")
    cond = nu.prepare(tok_ref.half(), pooled_ref.half(), img.half(), 16 * 16)
    got = nu.forward(_nhwc(lat).half(), nu.embeddings(t, cond))
    _close(got, ref, 0.04)
    z = torch.randn(1, 4, 8, 8, device="cuda")
    _close(NativeMoVQDecoder(movq.state_dict(), MoVQConfig.tiny(), dt, "cuda").forward(_nhwc(z).half()), _nhwc(movq(z)), 0.04)


@pytest.mark.gpu
def test_native_pipeline_matches_eager_and_is_reproducible():
    cfg = Kandinsky2Config.tiny()
    eager = Kandinsky2Pipeline(cfg, backend="eager", device="cuda", dtype=torch.float32).generate("arbius test cat", 32, 32, 1337)
    nat = Kandinsky2Pipeline(cfg, backend="native", device="cuda", dtype=torch.float16)
    a, b = nat.generate("arbius test cat", 32, 32, 1337), nat.generate("arbius test cat", 32, 32, 1337)
    assert (a == b).all() and a.shape == (32, 32, 3)
    diff = np.abs(a.astype(np.int32) - eager.astype(np.int32))
    assert diff.mean() < 3.0, (diff.mean(), diff.max())
