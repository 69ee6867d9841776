"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DT = [torch.float16, torch.bfloat16]


def _tol(dt):
    return (2e-2, 2e-2) if dt == torch.bfloat16 else (4e-3, 4e-3)


def _close(got, ref, dt, scale=1.0):
    rtol, atol = _tol(dt)
    got, ref = got.float(), ref.float()
    err = (got - ref).abs().max().item()
    bound = atol * scale + rtol * ref.abs().max().item()
    assert torch.isfinite(got).all(), "non-finite output"
    assert err <= bound, f"max abs err {err:.5f} > {bound:.5f} (ref max {ref.abs().max().item():.3f})"


@pytest.fixture(autouse=True)
def _seed():
    torch.manual_seed(1234)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("m,n,k,bn", [(128, 64, 64, 64), (256, 128, 128, 128), (8192, 320, 320, 0),
                                      (154, 320, 768, 0), (2048, 640, 2560, 0), (130, 1280, 1280, 32),
                                      (512, 1280, 320, 96), (1000, 256, 200, 160), (384, 512, 512, 256)])
def test_gemm_plain(dt, m, n, k, bn):
    from arbius_b200 import ops
    x = torch.randn(m, k, device="cuda", dtype=dt)
    w = torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5
    b = torch.randn(n, device="cuda", dtype=dt)
    y = ops.gemm(x, w, b, block_n=bn, splitk=1)
    ref = x.float() @ w.float().t() + b.float()
    _close(y, ref, dt)


@pytest.mark.parametrize("dt", DT)
def test_gemm_epilogues(dt):
    from arbius_b200 import ops
    m, n, k = 512, 640, 320
    x = torch.randn(m, k, device="cuda", dtype=dt)
    w = torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5
    b = torch.randn(n, device="cuda", dtype=dt)
    res = torch.randn(m, n, device="cuda", dtype=dt)
    rb = torch.randn(2, n, device="cuda", dtype=dt)
    base = x.float() @ w.float().t() + b.float()
    _close(ops.gemm(x, w, b, act="silu", splitk=1), F.silu(base), dt)
    _close(ops.gemm(x, w, b, act="gelu", splitk=1), F.gelu(base), dt)
    _close(ops.gemm(x, w, b, act="quick_gelu", splitk=1), base * torch.sigmoid(1.702 * base), dt)
    _close(ops.gemm(x, w, b, residual=res, splitk=1), base + res.float(), dt)
    _close(ops.gemm(x, w, b, rowbias=rb, rows_per_group=256, splitk=1),
           base + rb.float().repeat_interleave(256, 0), dt)
    _close(ops.gemm(x, w, None, out_scale=0.5, splitk=1), 0.5 * (x.float() @ w.float().t()), dt)


@pytest.mark.parametrize("dt", DT)
def test_gemm_geglu(dt):
    from arbius_b200 import ops
    m, k, inner = 256, 320, 1280
    x = torch.randn(m, k, device="cuda", dtype=dt)
    w = torch.randn(2 * inner, k, device="cuda", dtype=dt) / k ** 0.5
    b = torch.randn(2 * inner, device="cuda", dtype=dt)
    wp, bp = ops.pack_geglu_weight(w, b)
    y = ops.gemm(x, wp, bp, geglu=True, splitk=1)
    h = x.float() @ w.float().t() + b.float()
    ref = h[:, :inner] * F.gelu(h[:, inner:])
    assert y.shape == (m, inner)
    _close(y, ref, dt, scale=4)


@pytest.mark.parametrize("dt", DT)
def test_gemm_concat_and_splitk(dt):
    from arbius_b200 import ops
    m, k1, k2, n = 128, 1280, 640, 1280
    x1 = torch.randn(m, k1, device="cuda", dtype=dt)
    x2 = torch.randn(m, k2, device="cuda", dtype=dt)
    w = torch.randn(n, k1 + k2, device="cuda", dtype=dt) / (k1 + k2) ** 0.5
    b = torch.randn(n, device="cuda", dtype=dt)
    ref = torch.cat([x1, x2], 1).float() @ w.float().t() + b.float()
    _close(ops.gemm(x1, w, b, x2=x2, splitk=1), ref, dt)
    y4 = ops.gemm(x1, w, b, x2=x2, splitk=4, act="silu")
    _close(y4, F.silu(ref), dt)
    # fixed split => bit-identical on repeat
    assert torch.equal(y4, ops.gemm(x1, w, b, x2=x2, splitk=4, act="silu"))


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("b,h,w,cin,cout", [(2, 64, 64, 320, 320), (2, 32, 32, 640, 640), (2, 16, 16, 1280, 1280),
                                             (2, 8, 8, 1280, 1280), (1, 24, 40, 64, 128), (3, 9, 16, 128, 64),
                                             (1, 128, 128, 128, 128), (1, 96, 96, 64, 96)])
def test_conv3x3(dt, b, h, w, cin, cout):
    from arbius_b200 import ops
    x = torch.randn(b, h, w, cin, device="cuda", dtype=dt)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=dt) / (9 * cin) ** 0.5
    bias = torch.randn(cout, device="cuda", dtype=dt)
    y = ops.conv3x3(x, ops.pack_conv_weight(wt), bias)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), bias.float(), padding=1).permute(0, 2, 3, 1)
    _close(y, ref, dt)


@pytest.mark.parametrize("dt", DT)
def test_conv3x3_fused_epilogue_concat_splitk(dt):
    from arbius_b200 import ops
    b, h, w, c1, c2, cout = 2, 16, 16, 640, 320, 640
    x1 = torch.randn(b, h, w, c1, device="cuda", dtype=dt)
    x2 = torch.randn(b, h, w, c2, device="cuda", dtype=dt)
    wt = torch.randn(cout, c1 + c2, 3, 3, device="cuda", dtype=dt) / (9 * (c1 + c2)) ** 0.5
    bias = torch.randn(cout, device="cuda", dtype=dt)
    temb = torch.randn(b, cout, device="cuda", dtype=dt)
    res = torch.randn(b, h, w, cout, device="cuda", dtype=dt)
    xin = torch.cat([x1, x2], -1).float().permute(0, 3, 1, 2)
    ref = F.conv2d(xin, wt.float(), bias.float(), padding=1).permute(0, 2, 3, 1)
    ref = ref + temb.float()[:, None, None, :] + res.float()
    wp = ops.pack_conv_weight(wt)
    _close(ops.conv3x3(x1, wp, bias, x2=x2, rowbias=temb, residual=res, splitk=1), ref, dt)
    _close(ops.conv3x3(x1, wp, bias, x2=x2, rowbias=temb, residual=res, splitk=3), ref, dt)


@pytest.mark.parametrize("dt", DT)
def test_conv_strided_and_direct(dt):
    from arbius_b200 import ops
    x = torch.randn(2, 32, 32, 320, device="cuda", dtype=dt)
    wt = torch.randn(320, 320, 3, 3, device="cuda", dtype=dt) / (9 * 320) ** 0.5
    bias = torch.randn(320, device="cuda", dtype=dt)
    y = ops.conv3x3_strided(x, ops.pack_conv_weight(wt), bias, stride=2)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wt.float(), bias.float(), padding=1, stride=2).permute(0, 2, 3, 1)
    _close(y, ref, dt)
    # conv_in: 4 -> 320
    x4 = torch.randn(2, 64, 64, 4, device="cuda", dtype=dt)
    w4 = torch.randn(320, 4, 3, 3, device="cuda", dtype=dt) / 6
    y = ops.conv_direct(x4, w4.permute(0, 2, 3, 1).contiguous(), bias)
    ref = F.conv2d(x4.float().permute(0, 3, 1, 2), w4.float(), bias.float(), padding=1).permute(0, 2, 3, 1)
    _close(y, ref, dt)
    # conv_out: 320 -> 4
    wo = torch.randn(4, 320, 3, 3, device="cuda", dtype=dt) / (9 * 320) ** 0.5
    bo = torch.randn(4, device="cuda", dtype=dt)
    y = ops.conv_direct(x, wo.permute(0, 2, 3, 1).contiguous(), bo)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wo.float(), bo.float(), padding=1).permute(0, 2, 3, 1)
    _close(y, ref, dt)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("b,h,w,cin,cout,kw", [(2, 32, 32, 640, 640, {}), (2, 8, 8, 1280, 1280, {}), (1, 12, 20, 64, 128, {}),
                                                (3, 5, 9, 128, 64, {}), (2, 16, 16, 1280, 1280, dict(splitk=3, block_n=128)),
                                                (1, 64, 64, 256, 256, dict(block_n=256)), (1, 128, 128, 128, 128, {})])
def test_conv3x3_with_folded_upsample_equals_the_unfused_bytes(dt, b, h, w, cin, cout, kw):
    """K11: the nearest-2x upsample lives in the conv's TMA coordinates (one tile = one parity class of the output)."""
    from arbius_b200 import ops
    x = torch.randn(b, h, w, cin, device="cuda", dtype=dt)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=dt) / (9 * cin) ** 0.5
    bias = torch.randn(cout, device="cuda", dtype=dt)
    wp = ops.pack_conv_weight(wt)
    fused = ops.conv3x3(x, wp, bias, upsample=True, **kw)
    plain = ops.conv3x3(ops.upsample_nearest2x(x), wp, bias, **kw)
    assert fused.shape == (b, 2 * h, 2 * w, cout)
    assert torch.equal(fused, plain)
    up = F.interpolate(x.float().permute(0, 3, 1, 2), scale_factor=2.0, mode="nearest")
    ref = F.conv2d(up, wt.float(), bias.float(), padding=1).permute(0, 2, 3, 1)
    _close(fused, ref, dt)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("b,h,w,cin,cout,pad_mode", [(2, 64, 64, 320, 320, 0), (2, 16, 16, 1280, 1280, 0), (1, 33, 47, 64, 128, 0),
                                                      (1, 64, 64, 128, 128, 1), (2, 30, 22, 128, 64, 1), (1, 512, 512, 128, 128, 1)])
def test_conv3x3_stride2_through_tma_strides_equals_the_gathered_bytes(dt, b, h, w, cin, cout, pad_mode):
    """K11: stride 2 is the traversal stride of the A tensor map (no im2col copy); both paddings of the models."""
    from arbius_b200 import ops
    x = torch.randn(b, h, w, cin, device="cuda", dtype=dt)
    wt = torch.randn(cout, cin, 3, 3, device="cuda", dtype=dt) / (9 * cin) ** 0.5
    bias = torch.randn(cout, device="cuda", dtype=dt)
    wp = ops.pack_conv_weight(wt)
    fused = ops.conv3x3_strided(x, wp, bias, stride=2, pad_mode=pad_mode)
    gathered = ops.conv3x3_strided_gather(x, wp, bias, stride=2, pad_mode=pad_mode)
    assert fused.shape == gathered.shape
    assert torch.equal(fused, gathered)
    xin = x.float().permute(0, 3, 1, 2)
    if pad_mode == 1:
        xin = F.pad(xin, (0, 1, 0, 1))
    ref = F.conv2d(xin, wt.float(), bias.float(), padding=1 if pad_mode == 0 else 0, stride=2).permute(0, 2, 3, 1)
    _close(fused, ref, dt)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("m,n,k,kw", [(4096, 320, 320, {}), (2560 + 77, 640, 640, dict(act="silu")), (3000, 960, 320, dict(res=True)),
                                       (4096, 2560, 320, dict(geglu=True)), (2048 + 130, 5120, 640, dict(geglu=True, res=True)),
                                       (1024, 1280, 1280, dict(act="gelu")), (2304, 96, 128, {}), (5000, 352, 192, dict(act="silu", res=True))])
def test_two_sm_gemm_returns_the_one_sm_bytes(dt, m, n, k, kw):
    """csrc/gemm_2sm.cu (tcgen05.mma.cta_group::2, M-persistent, TMA-store epilogue) against csrc/gemm_tc.cu: the K order
    of every output element is the same, so the bytes must be; and both against fp32 PyTorch."""
    import ctypes as C
    from arbius_b200 import ops
    from arbius_b200.ops import native
    lib = native.load()
    lib.ab_gemm_set_2sm.argtypes = [C.c_int, C.c_int]
    lib.ab_gemm_set_2sm.restype = None
    geglu = kw.get("geglu", False)
    nout = n // 2 if geglu else n
    x = torch.randn(m, k, device="cuda", dtype=dt)
    w = torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5
    b = torch.randn(n, device="cuda", dtype=dt)
    res = torch.randn(m, nout, device="cuda", dtype=dt) if kw.get("res") else None
    args = dict(geglu=geglu, act=kw.get("act"), residual=res, splitk=1)
    try:
        lib.ab_gemm_set_2sm(0, -1)
        one = ops.gemm(x, w, b, **args)
        lib.ab_gemm_set_2sm(1, 256)
        two = ops.gemm(x, w, b, **args)
    finally:
        lib.ab_gemm_set_2sm(-1, -1)
    assert torch.equal(one, two)
    y = x.float() @ w.float().t() + b.float()
    if not geglu:
        if kw.get("act") == "silu":
            y = F.silu(y)
        elif kw.get("act") == "gelu":
            y = F.gelu(y)
        if res is not None:
            y = y + res.float()
        _close(two, y, dt)


@pytest.mark.parametrize("dt", DT)
def test_two_sm_conv_returns_the_one_sm_bytes(dt):
    """The pair tiles of csrc/gemm_2sm.cu on the implicit-GEMM convs: 3x3 with time-embedding row bias, residual, virtual
    concat; an odd number of 128-pixel tiles; tiles that span images; the temporal 3x1 conv."""
    import ctypes as C
    from arbius_b200 import ops
    from arbius_b200.ops import native
    lib = native.load()
    lib.ab_gemm_set_2sm.argtypes = [C.c_int, C.c_int]
    lib.ab_gemm_set_2sm.restype = None
    cases = []
    for (b, h, w, c1, c2, cout, act) in [(2, 64, 64, 320, 0, 320, "silu"), (3, 24, 40, 128, 64, 192, None), (8, 8, 8, 256, 0, 256, None),
                                         (5, 16, 16, 128, 128, 160, "silu")]:
        x1 = torch.randn(b, h, w, c1, device="cuda", dtype=dt)
        x2 = torch.randn(b, h, w, c2, device="cuda", dtype=dt) if c2 else None
        wp = ops.pack_conv_weight(torch.randn(cout, c1 + c2, 3, 3, device="cuda", dtype=dt) / (9 * (c1 + c2)) ** 0.5)
        bias = torch.randn(cout, device="cuda", dtype=dt)
        temb = torch.randn(b, cout, device="cuda", dtype=dt)
        res = torch.randn(b, h, w, cout, device="cuda", dtype=dt)
        cases.append(lambda x1=x1, x2=x2, wp=wp, bias=bias, temb=temb, res=res, act=act:
                     ops.conv3x3(x1, wp, bias, x2=x2, rowbias=temb, residual=res, act=act, splitk=1))
    xt = torch.randn(2, 12, 600, 128, device="cuda", dtype=dt)
    wt = ops.pack_temporal_conv_weight(torch.randn(192, 128, 3, 1, 1, device="cuda", dtype=dt) / (3 * 128) ** 0.5)
    bt = torch.randn(192, device="cuda", dtype=dt)
    cases.append(lambda: ops.conv_temporal(xt, wt, bt, splitk=1))
    for fn in cases:
        try:
            lib.ab_gemm_set_2sm(0, -1)
            one = fn()
            lib.ab_gemm_set_2sm(1, 256)
            two = fn()
        finally:
            lib.ab_gemm_set_2sm(-1, -1)
        assert torch.equal(one, two)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("n,hw,c1,c2,act", [(2, 4096, 320, 0, "silu"), (2, 1024, 1280, 640, "silu"),
                                            (2, 64, 1280, 1280, None), (1, 16384, 128, 0, "silu"),
                                            (3, 100, 64, 0, None), (2, 256, 1920, 0, "silu")])
def test_groupnorm(dt, n, hw, c1, c2, act):
    from arbius_b200 import ops
    x = torch.randn(n, hw, c1, device="cuda", dtype=dt) * 2 + 0.5
    x2 = torch.randn(n, hw, c2, device="cuda", dtype=dt) - 1 if c2 else None
    c = c1 + c2
    g = torch.randn(c, device="cuda", dtype=dt)
    b = torch.randn(c, device="cuda", dtype=dt)
    y = ops.groupnorm(x, g, b, 32, 1e-5, act, x2=x2)
    xin = torch.cat([x, x2], -1) if c2 else x
    ref = F.group_norm(xin.float().permute(0, 2, 1), 32, g.float(), b.float(), 1e-5).permute(0, 2, 1)
    if act == "silu":
        ref = F.silu(ref)
    _close(y, ref, dt)
    assert torch.equal(y, ops.groupnorm(x, g, b, 32, 1e-5, act, x2=x2)), "GroupNorm must be bit-reproducible"


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("rows,cols", [(154, 768), (8192, 320), (77, 1024), (33, 2048), (512, 1280)])
def test_layernorm(dt, rows, cols):
    from arbius_b200 import ops
    x = torch.randn(rows, cols, device="cuda", dtype=dt) * 3 + 1
    g = torch.randn(cols, device="cuda", dtype=dt)
    b = torch.randn(cols, device="cuda", dtype=dt)
    _close(ops.layernorm(x, g, b, 1e-5), F.layer_norm(x.float(), (cols,), g.float(), b.float(), 1e-5), dt)


def test_philox_matches_cpu_twin_and_is_shard_invariant():
    from arbius_b200 import ops
    full = ops.philox_normal((4, 64, 64), seed=1337, stream_id=3)
    cpu = ops.philox_normal_cpu(4 * 64 * 64, 1337, 3).reshape(4, 64, 64)
    assert (full.cpu() - cpu).abs().max().item() < 1e-5
    flat = full.reshape(-1)
    for off, n in [(0, 5), (3, 1001), (4096, 4096), (16383, 1)]:
        part = ops.philox_normal((n,), seed=1337, stream_id=3, offset=off)
        assert torch.equal(part, flat[off:off + n])
    assert abs(flat.mean().item()) < 0.05 and abs(flat.std().item() - 1) < 0.05


@pytest.mark.parametrize("dt", DT)
def test_elementwise(dt):
    from arbius_b200 import ops
    ec = torch.randn(2 * 64 * 64 * 4, device="cuda", dtype=dt)
    eu = torch.randn_like(ec)
    x = torch.randn(ec.numel(), device="cuda")
    nz = torch.randn_like(x)
    xo = torch.empty_like(x)
    xm = torch.empty_like(ec)
    ops.cfg_step(ec, eu, x, g=7.5, ca=0.9, cb=-0.3, noise=nz, cn=0.1, x_out=xo, xm_out=xm)
    eps = eu.float() + 7.5 * (ec.float() - eu.float())
    ref = 0.9 * x - 0.3 * eps + 0.1 * nz
    assert (xo - ref).abs().max().item() < 1e-4
    _close(xm, ref, dt)
    img = torch.randn(1, 32, 32, 3, device="cuda", dtype=dt)
    u8 = ops.finalize_u8(img)
    refu = ((img.float() / 2 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8)
    assert (u8.int() - refu.int()).abs().max().item() <= 0
    t = torch.randn(2, 8, 8, 64, device="cuda", dtype=dt)
    up = ops.upsample_nearest2x(t)
    assert torch.equal(up, t.repeat_interleave(2, 1).repeat_interleave(2, 2))
    rs = ops.resize_bilinear(t, 20, 12)
    refr = F.interpolate(t.float().permute(0, 3, 1, 2), size=(20, 12), mode="bilinear", align_corners=False)
    _close(rs, refr.permute(0, 2, 3, 1), dt)
    a, b = torch.randn(4096, device="cuda", dtype=dt), torch.randn(4096, device="cuda", dtype=dt)
    _close(ops.binary(a, b, "add", act="silu"), F.silu(a.float() + b.float()), dt)
    ts = torch.tensor([999.0, 500.0], device="cuda")
    te = ops.timestep_embedding(ts, 320, dt)
    half = 160
    freqs = torch.exp(-torch.log(torch.tensor(10000.0)) * torch.arange(half, device="cuda") / half)
    args = ts[:, None] * freqs[None]
    _close(te, torch.cat([args.cos(), args.sin()], -1), dt)


def _attn_ref(q, k, v, scale=None, causal=False):
    qf, kf, vf = (t.float().permute(0, 2, 1, 3) for t in (q, k, v))
    o = F.scaled_dot_product_attention(qf, kf, vf, is_causal=causal, scale=scale)
    return o.permute(0, 2, 1, 3)


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("b,h,lq,lk,d,causal", [(2, 8, 4096, 4096, 40, False), (2, 8, 1024, 1024, 80, False),
                                                 (2, 8, 256, 256, 160, False), (2, 8, 64, 64, 160, False),
                                                 (2, 8, 4096, 77, 40, False), (2, 8, 1024, 77, 80, False),
                                                 (2, 12, 77, 77, 64, True), (1, 5, 9216, 9216, 64, False),
                                                 (3, 4, 300, 500, 64, False), (2, 2, 200, 200, 128, True)])
def test_flash_attn(dt, b, h, lq, lk, d, causal):
    from arbius_b200 import ops
    q = torch.randn(b, lq, h, d, device="cuda", dtype=dt)
    k = torch.randn(b, lk, h, d, device="cuda", dtype=dt)
    v = torch.randn(b, lk, h, d, device="cuda", dtype=dt)
    o = ops.flash_attn(q, k, v, causal=causal)
    _close(o, _attn_ref(q, k, v, causal=causal), dt)
    assert torch.equal(o, ops.flash_attn(q, k, v, causal=causal))


@pytest.mark.parametrize("dt", DT)
def test_flash_attn_fused_qkv_views_and_large_logits(dt):
    from arbius_b200 import ops
    b, l, h, d = 2, 640, 8, 40
    qkv = torch.randn(b, l, 3 * h * d, device="cuda", dtype=dt)
    q, k, v = (qkv[..., i * h * d:(i + 1) * h * d].view(b, l, h, d) for i in range(3))
    _close(ops.flash_attn(q, k, v), _attn_ref(q, k, v), dt)
    # logits with a huge dynamic range exercise the lazy-rescale path
    q2 = q * 6
    _close(ops.flash_attn(q2, k, v), _attn_ref(q2, k, v), dt)


@pytest.mark.parametrize("dt", DT)
def test_gemm_two_tiles_per_cta_and_tma_store_edges(dt):
    """Large BN=256 problems take the two-M-tiles-per-CTA kernel; odd widths and sliced outputs go through the
    TMA-store epilogue (clipped boxes) or its direct-store fallback."""
    from arbius_b200 import ops
    x = torch.randn(16384, 1536, device="cuda", dtype=dt)
    w = torch.randn(2048, 1536, device="cuda", dtype=dt) / 1536 ** 0.5
    b = torch.randn(2048, device="cuda", dtype=dt)
    res = torch.randn(16384, 2048, device="cuda", dtype=dt)
    y = ops.gemm(x, w, b, residual=res, act="silu", block_n=256, splitk=1)
    ref = F.silu(x.float() @ w.float().t() + b.float()) + res.float()
    _close(y, ref, dt)
    # conv through the same path: 16 images of 64x64, 256 -> 512 channels
    xi = torch.randn(16, 64, 64, 256, device="cuda", dtype=dt)
    wc = torch.randn(512, 256, 3, 3, device="cuda", dtype=dt) / (9 * 256) ** 0.5
    yc = ops.conv3x3(xi, ops.pack_conv_weight(wc), None, block_n=256, splitk=1)
    refc = F.conv2d(xi.float().permute(0, 3, 1, 2), wc.float(), padding=1).permute(0, 2, 3, 1)
    _close(yc, refc, dt)
    # N = 72 (last 32-column box clipped to 8 columns), M = 300 (last row box clipped), output a column slice
    x2 = torch.randn(300, 200, device="cuda", dtype=dt)
    w2 = torch.randn(72, 200, device="cuda", dtype=dt) / 200 ** 0.5
    wide = torch.zeros(300, 160, device="cuda", dtype=dt)
    ops.gemm(x2, w2, None, out=wide[:, 80:152], splitk=1)
    _close(wide[:, 80:152], x2.float() @ w2.float().t(), dt)
    assert (wide[:, :80] == 0).all() and (wide[:, 152:] == 0).all(), "TMA store wrote outside its column slice"


def test_groupnorm_fused_kernel_matches_two_kernel_path_bitwise():
    """ab_groupnorm (one cooperative launch when the rows fit shared memory) and ab_groupnorm_static (always the
    two-kernel path) share the canonical decomposition: same bits."""
    import ctypes as C
    from arbius_b200 import ops
    from arbius_b200.ops import native
    from arbius_b200.ops.native import ACT, GroupNormArgs
    lib = native.load()
    for (n, hw, c) in [(2, 4096, 320), (2, 1024, 640), (1, 256, 1280)]:
        x = torch.randn(n, hw, c, device="cuda", dtype=torch.float16) * 3 - 1
        g = torch.randn(c, device="cuda", dtype=torch.float16)
        b = torch.randn(c, device="cuda", dtype=torch.float16)
        y = ops.groupnorm(x, g, b, 32, 1e-5, "silu")
        y2 = torch.empty_like(x)
        a = GroupNormArgs()
        a.x, a.y, a.gamma, a.beta = x.data_ptr(), y2.data_ptr(), g.data_ptr(), b.data_ptr()
        a.n, a.hw, a.c1, a.c2, a.groups, a.eps, a.act, a.dtype = n, hw, c, 0, 32, 1e-5, ACT["silu"], 0
        lib.ab_groupnorm_static.restype = C.c_int
        lib.ab_groupnorm_static.argtypes = [C.POINTER(GroupNormArgs), C.c_void_p]
        native.check(lib.ab_groupnorm_static(C.byref(a), native.stream_ptr()), "ab_groupnorm_static")
        torch.cuda.synchronize()
        assert torch.equal(y, y2), (n, hw, c)

