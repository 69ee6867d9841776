"""Robust Video Matting: eager pipeline + video IO on CPU; matting kernels, native engine and pipeline vs eager on GPU."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from arbius_b200.codecs.mp4 import encode_mp4, probe_fps
from arbius_b200.models.rvm.pipeline import RvmPipeline, composite_reference, decode_video
from arbius_b200.models.rvm.ref import MattingNetwork, RvmConfig, init_rvm_


def _clip(T=3, H=64, W=96, seed=0):
    rng = np.random.RandomState(seed)
    base = rng.randint(0, 255, (H // 8, W // 8, 3)).astype(np.uint8)
    frames = []
    for t in range(T):
        f = np.kron(np.roll(base, t, axis=1), np.ones((8, 8, 1), dtype=np.uint8))
        frames.append(f)
    return np.stack(frames).astype(np.uint8)


def test_state_dict_layout_and_size():
    m = MattingNetwork()
    keys = set(m.state_dict())
    for k in ("backbone.features.0.0.weight", "aspp.aspp1.0.weight", "decoder.decode4.gru.ih.0.weight",
              "decoder.decode3.conv.0.weight", "decoder.decode0.conv.3.weight", "project_mat.conv.weight",
              "refiner.box_filter.weight", "refiner.conv.6.bias"):
        assert k in keys, k
    assert m.state_dict()["decoder.decode3.conv.0.weight"].shape == (80, 171, 3, 3)
    assert abs(sum(p.numel() for p in m.parameters()) / 1e6 - 3.75) < 0.05


def test_eager_pipeline_cpu_recurrent_and_deterministic():
    pipe = RvmPipeline(RvmConfig.tiny(), backend="eager", device="cpu", dtype=torch.float32)
    clip = _clip()
    a, b = pipe.matte(clip, "green-screen"), pipe.matte(clip, "green-screen")
    assert a.shape == clip.shape and a.dtype == np.uint8 and (a == b).all()
    # recurrent: the same picture at t=0 and as a later frame gives different mattes
    same = np.stack([clip[0], clip[0]])
    out = pipe.matte(same, "alpha-mask")
    assert (out[0] != out[1]).any()
    assert (out[..., 0] == out[..., 1]).all()                   # alpha mask is grey


def test_predict_reads_and_writes_mp4():
    pipe = RvmPipeline(RvmConfig.tiny(), backend="eager", device="cpu", dtype=torch.float32)
    clip = _clip(T=2)
    data = encode_mp4(clip, 12)
    frames, fps = decode_video(data)
    assert fps == 12 and probe_fps(data) == 12 and frames.shape == clip.shape
    assert np.abs(frames.astype(int) - clip.astype(int)).mean() < 6.0     # JPEG q90
    out = pipe.predict({"input_video": data, "output_type": ""})
    out2 = pipe.predict({"input_video": data, "output_type": "green-screen"})
    assert out == out2 and out[4:8] == b"ftyp"
    f2, fps2 = decode_video(out)
    assert f2.shape == clip.shape and fps2 == 12


def _close(got, ref, tol):
    got, ref = got.float(), ref.float()
    assert torch.isfinite(got).all()
    err = (got - ref).abs().max().item()
    assert err <= tol * max(1.0, ref.abs().max().item()), err


def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


@pytest.mark.gpu
@pytest.mark.parametrize("cs,cout,ks,stride,dil", [((3,), 16, 3, 2, 1), ((128, 40, 3), 80, 3, 1, 1), ((32, 3), 16, 3, 1, 1),
                                                   ((16,), 4, 1, 1, 1), ((8, 16), 16, 1, 1, 1), ((24,), 40, 3, 1, 2)])
def test_conv_generic_matches_torch(cs, cout, ks, stride, dil):
    from arbius_b200.ops import matting as M
    torch.manual_seed(0)
    n, h, w = 2, 19, 23
    xs = [torch.randn(n, sum(cs) if False else c, h, w, device="cuda") for c in cs]
    wt = torch.randn(cout, sum(cs), ks, ks, device="cuda") / (sum(cs) * ks * ks) ** 0.5
    b = torch.randn(cout, device="cuda") * 0.1
    ref = F.relu(F.conv2d(torch.cat(xs, 1).half().float(), wt.half().float(), b.half().float(), stride=stride,
                          padding=(ks // 2) * dil, dilation=dil))
    got = M.conv_generic([_nhwc(x).half() for x in xs], M.pack_generic_weight(wt, torch.float16, "cuda"), b.half(), cout, ks,
                         stride=stride, dilation=dil, act="relu")
    _close(got, _nhwc(ref), 4e-3)


@pytest.mark.gpu
def test_conv_generic_sliced_sources_gate_affine_and_gru():
    from arbius_b200.ops import matting as M
    torch.manual_seed(1)
    n, h, w, c = 1, 13, 17, 16
    wide = torch.randn(n, h, w, 2 * c, device="cuda").half()
    hprev = torch.randn(n, h, w, c, device="cuda").half()
    rz = torch.rand(n, h, w, 2 * c, device="cuda").half()
    wt = torch.randn(c, 2 * c, 3, 3, device="cuda") / (2 * c * 9) ** 0.5
    b = torch.randn(c, device="cuda") * 0.1
    xb = wide[..., c:]
    got = M.conv_generic([xb, hprev], M.pack_generic_weight(wt, torch.float16, "cuda"), b.half(), c, 3, act="tanh",
                         gates=[None, rz[..., :c]], gru=(rz[..., c:], hprev))
    inp = torch.cat([xb.float(), hprev.float() * rz[..., :c].float()], -1).permute(0, 3, 1, 2)
    cand = torch.tanh(F.conv2d(inp, wt.half().float(), b.half().float(), padding=1)).permute(0, 2, 3, 1)
    z = rz[..., c:].float()
    _close(got, (1 - z) * hprev.float() + z * cand, 4e-3)
    # input affine on a 3-channel source: padding stays zero in the normalised domain
    x = torch.rand(1, 9, 11, 3, device="cuda").half()
    sc, sh = torch.tensor([2.0, 3.0, 4.0], device="cuda"), torch.tensor([-1.0, 0.5, 0.25], device="cuda")
    w3 = torch.randn(16, 3, 3, 3, device="cuda") / 27 ** 0.5
    got = M.conv_generic([x], M.pack_generic_weight(w3, torch.float16, "cuda"), None, 16, 3, in_affine=(sc, sh))
    ref = F.conv2d((x.float() * sc + sh).permute(0, 3, 1, 2), w3.half().float(), padding=1)
    _close(got, _nhwc(ref), 4e-3)


@pytest.mark.gpu
def test_resize_pool_copy_kernels():
    from arbius_b200.ops import matting as M
    torch.manual_seed(2)
    u8 = torch.randint(0, 256, (1, 54, 96, 3), device="cuda", dtype=torch.uint8)
    got = M.resize_bilinear_ex(u8, 13, 24, 4.0, 4.0, dtype=torch.float16)
    ref = F.interpolate(u8.permute(0, 3, 1, 2).float() / 255, scale_factor=0.25, mode="bilinear", align_corners=False,
                        recompute_scale_factor=False)
    _close(got, _nhwc(ref), 2e-3)
    x = torch.randn(2, 9, 7, 24, device="cuda").half()
    got = M.resize_bilinear_ex(x, 17, 13, 0.5, 0.5)                # 2x upsample cropped to 17x13
    ref = F.interpolate(x.permute(0, 3, 1, 2).float(), scale_factor=2, mode="bilinear", align_corners=False)[:, :, :17, :13]
    _close(got, _nhwc(ref), 2e-3)
    x = torch.randn(2, 9, 7, 3, device="cuda").half()
    ref = F.avg_pool2d(x.permute(0, 3, 1, 2).float(), 2, 2, ceil_mode=True, count_include_pad=False)
    _close(M.avgpool2x_ceil(x), _nhwc(ref), 2e-3)
    src = torch.randn(1, 5, 6, 8, device="cuda").half()
    d1, d2 = torch.zeros(1, 5, 6, 16, device="cuda").half(), torch.zeros(1, 5, 6, 8, device="cuda").half()
    M.copy_cols(src, d1[..., 8:], d2)
    assert (d1[..., 8:] == src).all() and (d1[..., :8] == 0).all() and (d2 == src).all()


@pytest.mark.gpu
@pytest.mark.parametrize("ratio", [0.5, 1.0])
def test_native_engine_matches_eager_over_frames(ratio):
    from arbius_b200.models.rvm.native import NativeRvm
    cfg = RvmConfig(downsample_ratio=ratio)
    net = init_rvm_(MattingNetwork(cfg), 3).cuda()
    eng = NativeRvm(net.state_dict(), cfg, torch.float16, "cuda")
    clip = torch.from_numpy(_clip(T=3, H=72, W=104, seed=5)).cuda()
    rec = [None] * 4
    state = eng.new_state(1, 72, 104)
    for t in range(3):
        src = clip[t:t + 1]
        fgr_r, pha_r, *rec = net(src.permute(0, 3, 1, 2).float() / 255, *rec)
        fgr, pha, u8 = eng.forward(src, state, mode=0)
        _close(fgr, _nhwc(fgr_r), 0.03)
        _close(pha, _nhwc(pha_r), 0.03)
        for got, ref in zip(state, rec):
            _close(got, _nhwc(ref), 0.03)
        diff = (u8.int() - composite_reference(fgr_r, pha_r, 0).int()).abs()
        assert diff.float().mean() < 1.5 and diff.max() <= 12, (diff.float().mean(), diff.max())


@pytest.mark.gpu
def test_native_pipeline_graph_matches_eager_and_is_reproducible():
    cfg = RvmConfig.tiny()
    clip = _clip(T=4, H=64, W=96, seed=7)
    eager = RvmPipeline(cfg, backend="eager", device="cuda", dtype=torch.float32).matte(clip, "foreground-mask")
    nat = RvmPipeline(cfg, backend="native", device="cuda", dtype=torch.float16)
    a = nat.matte(clip, "foreground-mask")
    b = nat.matte(clip, "foreground-mask")                        # second call replays the captured graph from zero state
    assert (a == b).all() and a.shape == clip.shape
    nog = RvmPipeline(cfg, backend="native", device="cuda", dtype=torch.float16, use_graph=False).matte(clip, "foreground-mask")
    assert (a == nog).all()
    diff = np.abs(a.astype(np.int32) - eager.astype(np.int32))
    assert diff.mean() < 1.5, (diff.mean(), diff.max())
    assert nat.last_launches > 0
