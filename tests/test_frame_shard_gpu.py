"""Building blocks of the frame-sharded UNet3D, checked on ONE GPU: each sharded operator, fed the shards of a tensor one
"rank" after the other, must reproduce the undivided kernel's bytes."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from arbius_b200 import ops
    return ops


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("shape", [(2, 8, 36, 64), (1, 24, 9216 // 16, 320)])
def test_sharded_groupnorm_matches_undivided_bytes(world, shape):
    ops = _ops()
    from arbius_b200.ops.native import ACT, GroupNormArgs
    from arbius_b200.parallel.frame_shard import _lib
    b, f, hw, c = shape
    torch.manual_seed(0)
    x = (torch.randn(b, f * hw, c, device="cuda") * 2 + 0.3).half()
    gamma, beta = torch.randn(c, device="cuda").half(), torch.randn(c, device="cuda").half()
    ref = ops.groupnorm(x, gamma, beta, 32, 1e-5, "silu")
    lib = _lib()
    fl = f // world
    parts = [torch.zeros(b * 8 * 32 * 32 * 16, dtype=torch.uint8, device="cuda") for _ in range(world)]
    shards = [x.view(b, f, hw, c)[:, r * fl:(r + 1) * fl].contiguous().view(b, fl * hw, c) for r in range(world)]
    args = []
    for r in range(world):
        a = GroupNormArgs()
        a.x, a.gamma, a.beta = shards[r].data_ptr(), gamma.data_ptr(), beta.data_ptr()
        a.partials = ops.norm._gn_workspace(b, 32, x.device).data_ptr()
        a.n, a.hw, a.c1, a.c2, a.groups = b, fl * hw, c, 0, 32
        a.eps, a.act, a.dtype = 1e-5, ACT["silu"], ops.native.dtype_code(x.dtype)
        assert lib.ab_groupnorm_sharded_stats(C.byref(a), world, r, parts[r].data_ptr(), ops.native.stream_ptr()) == 0
        args.append(a)
    srcs = (C.c_void_p * 8)(*[p.data_ptr() for p in parts] + [None] * (8 - world))
    # rank r writes into the interior of a padded buffer, as the temporal conv wants it
    for r in range(world):
        pad = torch.full((b, fl + 2, hw, c), 7.0, dtype=x.dtype, device="cuda")
        args[r].y = pad[:, 1].data_ptr()
        assert lib.ab_groupnorm_sharded_apply(C.byref(args[r]), world, srcs, (fl + 2) * hw * c, ops.native.stream_ptr()) == 0
        got = pad[:, 1:fl + 1].reshape(b, fl * hw, c)
        want = ref.view(b, f, hw, c)[:, r * fl:(r + 1) * fl].reshape(b, fl * hw, c)
        assert torch.equal(got, want), f"rank {r}"
        assert (pad[:, 0] == 7).all() and (pad[:, -1] == 7).all()


@pytest.mark.parametrize("world", [2, 4])
def test_halo_temporal_conv_matches_undivided_bytes(world):
    ops = _ops()
    b, f, hw, c, n = 2, 8, 160, 128, 192
    torch.manual_seed(1)
    x = torch.randn(b, f, hw, c, device="cuda").half()
    w = ops.pack_temporal_conv_weight((torch.randn(n, c, 3, 1, 1, device="cuda") * 0.05).half())
    bias = torch.randn(n, device="cuda").half()
    res = torch.randn(b, f, hw, n, device="cuda").half()
    ref = ops.conv_temporal(x, w, bias, residual=res)
    fl = f // world
    zero = torch.zeros(b, 1, hw, c, dtype=x.dtype, device="cuda")
    with ops.batch_scale(world):
        for r in range(world):
            prev = x[:, r * fl - 1:r * fl] if r > 0 else zero
            nxt = x[:, (r + 1) * fl:(r + 1) * fl + 1] if r < world - 1 else zero
            pad = torch.cat([prev, x[:, r * fl:(r + 1) * fl], nxt], 1).contiguous()
            got = ops.conv_temporal(pad, w, bias, residual=res[:, r * fl:(r + 1) * fl].contiguous(), halo=True)
            assert torch.equal(got, ref[:, r * fl:(r + 1) * fl]), f"rank {r}"


@pytest.mark.parametrize("world,frames", [(2, 24), (4, 24), (8, 24), (2, 8), (2, 16)])
def test_multi_source_pixel_band_temporal_attention_matches_undivided_bytes(world, frames):
    ops = _ops()
    from arbius_b200.ops.temporal import temporal_attn, temporal_attn_ptrs
    b, hw, heads, d = 2, 203, 5, 64
    c = heads * d
    torch.manual_seed(2)
    qkv = torch.randn(b, frames, hw, 3 * c, device="cuda").half()
    q, k, v = (qkv[..., i * c:(i + 1) * c] for i in range(3))
    ref = temporal_attn(q, k, v, heads)
    fl = frames // world
    srcs = [qkv[:, r * fl:(r + 1) * fl].contiguous() for r in range(world)]
    outs = [torch.zeros(b, fl, hw, c, dtype=qkv.dtype, device="cuda") for _ in range(world)]
    qp = [s.data_ptr() for s in srcs]
    kp = [p + c * 2 for p in qp]
    vp = [p + 2 * c * 2 for p in qp]
    op = [o.data_ptr() for o in outs]
    px = 128 // frames
    tiles = (hw + px - 1) // px
    per = (tiles + world - 1) // world
    for r in range(world):                                   # one launch per "rank": its pixel band, all sources
        p0, p1 = min(hw, r * per * px), min(hw, (r + 1) * per * px)
        if p1 > p0:
            temporal_attn_ptrs(qp, kp, vp, op, b, frames, hw, heads, d, 3 * c, c, qkv.dtype, pix_begin=p0, pix_count=p1 - p0)
    got = torch.cat(outs, 1)
    assert torch.equal(got, ref)


def test_peer_copy2d_strided_and_zero_fill():
    ops = _ops()
    from arbius_b200.parallel.frame_shard import _lib
    lib = _lib()
    src = torch.arange(4 * 6 * 64, dtype=torch.int16, device="cuda").view(4, 6, 64)
    dst = torch.full((4, 3, 64), -1, dtype=torch.int16, device="cuda")
    assert lib.ab_peer_copy2d(dst[:, 1].data_ptr(), 3 * 128, src[:, 4].data_ptr(), 6 * 128, 128, 4, ops.native.stream_ptr()) == 0
    assert torch.equal(dst[:, 1], src[:, 4]) and (dst[:, 0] == -1).all() and (dst[:, 2] == -1).all()
    assert lib.ab_peer_copy2d(dst[:, 2].data_ptr(), 3 * 128, None, 0, 128, 4, ops.native.stream_ptr()) == 0
    assert (dst[:, 2] == 0).all() and torch.equal(dst[:, 1], src[:, 4])


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("concat", [False, True])
def test_halo_conv3x3_row_bands_match_undivided_bytes(world, concat):
    """Patch parallelism's conv: every rank holds a row band of the image plus one halo row above and below."""
    ops = _ops()
    b, h, wd, c, n = 2, 32, 32, 128, 192
    torch.manual_seed(3)
    x = torch.randn(b, h, wd, c, device="cuda").half()
    x2 = torch.randn(b, h, wd, 64, device="cuda").half() if concat else None
    cin = c + (64 if concat else 0)
    w = ops.pack_conv_weight((torch.randn(n, cin, 3, 3, device="cuda") * 0.03).half())
    bias = torch.randn(n, device="cuda").half()
    ref = ops.conv3x3(x, w, bias, x2=x2, act="silu")
    hl = h // world

    def band(t, r):
        z = torch.zeros(b, 1, wd, t.shape[-1], dtype=t.dtype, device="cuda")
        top = t[:, r * hl - 1:r * hl] if r > 0 else z
        bot = t[:, (r + 1) * hl:(r + 1) * hl + 1] if r < world - 1 else z
        return torch.cat([top, t[:, r * hl:(r + 1) * hl], bot], 1).contiguous()
    with ops.batch_scale(world):
        for r in range(world):
            got = ops.conv3x3(band(x, r), w, bias, x2=band(x2, r) if concat else None, act="silu", halo=True)
            assert torch.equal(got, ref[:, r * hl:(r + 1) * hl]), f"rank {r}"
