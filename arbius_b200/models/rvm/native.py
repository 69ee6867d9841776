"""Native RVM engine: MobileNetV3 encoder 1x1 convs on the tcgen05 GEMM, depthwise/SE/GRU/guided-filter kernels
from csrc/elementwise2.cu and csrc/matting.cu. Channels-last, BatchNorm folded into the conv weights at load.

Weights come from a ``MattingNetwork`` state dict (upstream key layout, see ref.py). One call = one frame; the four
recurrent states live in fixed buffers that are updated in place, so a frame is CUDA-graph capturable.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ... import ops
from ...ops import extra as X
from ...ops import matting as M
from .ref import IMAGENET_MEAN, IMAGENET_STD, RvmConfig

# MobileNetV3-Large (dilated): kernel, expanded, out, squeeze-excite, activation, stride, dilation
MBV3_LARGE = (
    (3, 16, 16, False, "relu", 1, 1), (3, 64, 24, False, "relu", 2, 1), (3, 72, 24, False, "relu", 1, 1),
    (5, 72, 40, True, "relu", 2, 1), (5, 120, 40, True, "relu", 1, 1), (5, 120, 40, True, "relu", 1, 1),
    (3, 240, 80, False, "hardswish", 2, 1), (3, 200, 80, False, "hardswish", 1, 1), (3, 184, 80, False, "hardswish", 1, 1),
    (3, 184, 80, False, "hardswish", 1, 1), (3, 480, 112, True, "hardswish", 1, 1), (3, 672, 112, True, "hardswish", 1, 1),
    (5, 672, 160, True, "hardswish", 1, 2), (5, 960, 160, True, "hardswish", 1, 2), (5, 960, 160, True, "hardswish", 1, 2),
)
FEATURE_TAPS = {1: "f1", 3: "f2", 6: "f3"}      # after features[i]


def _fold(sd: Dict[str, torch.Tensor], conv: str, bn: Optional[str], eps: float):
    w = sd[conv + ".weight"].float()
    b = sd.get(conv + ".bias")
    b = b.float() if b is not None else torch.zeros(w.shape[0])
    if bn is not None:
        g = sd[bn + ".weight"].float() / torch.sqrt(sd[bn + ".running_var"].float() + eps)
        w = w * g.view(-1, 1, 1, 1)
        b = (b - sd[bn + ".running_mean"].float()) * g + sd[bn + ".bias"].float()
    return w, b


class _Lin:
    """1x1 conv as a GEMM: weight [Cout, Cin]"""

    def __init__(self, w, b, dtype, device):
        self.w = w.reshape(w.shape[0], -1).to(device=device, dtype=dtype).contiguous()
        self.b = b.to(device=device, dtype=dtype).contiguous()

    def __call__(self, x, act=None, residual=None):
        return ops.gemm(x, self.w, self.b, act=act, residual=residual)


class _Gen:
    """generic direct conv (csrc/matting.cu)"""

    def __init__(self, w, b, dtype, device):
        self.cout, self.ks = w.shape[0], w.shape[2]
        self.w = M.pack_generic_weight(w, dtype, device)
        self.b = b.to(device=device, dtype=dtype).contiguous()

    def __call__(self, srcs, **kw):
        return M.conv_generic(srcs, self.w, self.b, self.cout, self.ks, **kw)


class _Gru:
    def __init__(self, sd, prefix, dtype, device):
        self.ih = _Gen(*_fold(sd, prefix + ".ih.0", None, 0.0), dtype, device)
        self.hh = _Gen(*_fold(sd, prefix + ".hh.0", None, 0.0), dtype, device)
        self.c = self.hh.cout

    def __call__(self, xfull: torch.Tensor, h: torch.Tensor) -> None:
        """xfull [N,H,W,2c]: its upper channel half is replaced by the new hidden state, h is updated in place."""
        c = self.c
        b = xfull[..., c:]
        rz = self.ih([b, h], act="sigmoid")
        hn = self.hh([b, h], act="tanh", gates=[None, rz[..., :c]], gru=(rz[..., c:], h))
        M.copy_cols(hn, b, h)


class NativeRvm:
    def __init__(self, sd: Dict[str, torch.Tensor], cfg: RvmConfig, dtype=torch.float16, device="cuda"):
        self.cfg, self.dtype, self.device = cfg, dtype, torch.device(device)
        sd = {k: v.detach().cpu() for k, v in sd.items()}
        dt, dev = dtype, self.device
        f = "backbone.features."
        self.stem = _Gen(*_fold(sd, f + "0.0", f + "0.1", 1e-3), dt, dev)
        std = torch.tensor(IMAGENET_STD)
        self.in_affine = ((1.0 / std).to(dev), (-torch.tensor(IMAGENET_MEAN) / std).to(dev))
        self.blocks: List[dict] = []
        cin = 16
        for i, (k, exp, cout, se, act, stride, dil) in enumerate(MBV3_LARGE, start=1):
            p = f"{f}{i}.block."
            j = 0
            blk = {"k": k, "act": act, "stride": stride, "dil": dil, "res": stride == 1 and cin == cout and not (i == 13)}
            if exp != cin:
                blk["expand"] = _Lin(*_fold(sd, f"{p}{j}.0", f"{p}{j}.1", 1e-3), dt, dev)
                j += 1
            w, b = _fold(sd, f"{p}{j}.0", f"{p}{j}.1", 1e-3)
            blk["dw_w"] = w.permute(2, 3, 0, 1).reshape(k * k, exp).to(device=dev, dtype=dt).contiguous()
            blk["dw_b"] = b.to(device=dev, dtype=dt).contiguous()
            j += 1
            if se:
                blk["fc1"] = _Lin(*_fold(sd, f"{p}{j}.fc1", None, 0.0), dt, dev)
                blk["fc2"] = _Lin(*_fold(sd, f"{p}{j}.fc2", None, 0.0), dt, dev)
                j += 1
            blk["project"] = _Lin(*_fold(sd, f"{p}{j}.0", f"{p}{j}.1", 1e-3), dt, dev)
            self.blocks.append(blk)
            cin = cout
        self.last = _Lin(*_fold(sd, f + "16.0", f + "16.1", 1e-3), dt, dev)
        self.aspp1 = _Lin(*_fold(sd, "aspp.aspp1.0", "aspp.aspp1.1", 1e-5), dt, dev)
        self.aspp2 = _Lin(*_fold(sd, "aspp.aspp2.1", None, 0.0), dt, dev)
        d = "decoder."
        self.gru4 = _Gru(sd, d + "decode4.gru", dt, dev)
        self.up = []
        for name in ("decode3", "decode2", "decode1"):
            self.up.append((_Gen(*_fold(sd, f"{d}{name}.conv.0", f"{d}{name}.conv.1", 1e-5), dt, dev),
                            _Gru(sd, f"{d}{name}.gru", dt, dev)))
        self.out0 = _Gen(*_fold(sd, d + "decode0.conv.0", d + "decode0.conv.1", 1e-5), dt, dev)
        self.out1 = _Gen(*_fold(sd, d + "decode0.conv.3", d + "decode0.conv.4", 1e-5), dt, dev)
        self.project = _Gen(*_fold(sd, "project_mat.conv", None, 0.0), dt, dev)
        self.ref1 = _Gen(*_fold(sd, "refiner.conv.0", "refiner.conv.1", 1e-5), dt, dev)
        self.ref2 = _Gen(*_fold(sd, "refiner.conv.3", "refiner.conv.4", 1e-5), dt, dev)
        self.ref3 = _Gen(*_fold(sd, "refiner.conv.6", None, 0.0), dt, dev)
        self.state_channels = (16, 20, 40, 64)

    # -------------------------------------------------------------------------------------------------
    def working_size(self, H: int, W: int):
        r = self.cfg.downsample_ratio
        return (H, W) if r == 1 else (int(H * r), int(W * r))

    def new_state(self, n: int, H: int, W: int) -> List[torch.Tensor]:
        h, w = self.working_size(H, W)
        out = []
        for c in self.state_channels:
            h, w = (h + 1) // 2, (w + 1) // 2
            out.append(torch.zeros((n, h, w, c), dtype=self.dtype, device=self.device))
        return out

    def _encoder(self, src_sm: torch.Tensor):
        x = self.stem([src_sm], stride=2, act="hardswish", in_affine=self.in_affine)
        feats = {}
        for i, blk in enumerate(self.blocks, start=1):
            inp = x
            if "expand" in blk:
                x = blk["expand"](x, act=blk["act"])
            x = X.dwconv(x, blk["dw_w"], blk["dw_b"], blk["k"], blk["stride"], blk["dil"], act=blk["act"])
            if "fc1" in blk:
                s = X.global_avgpool(x)
                s = blk["fc2"](blk["fc1"](s, act="relu"), act="hardsigmoid")
                x = X.channel_scale(x, s)
            x = blk["project"](x, residual=inp if blk["res"] else None)
            if i in FEATURE_TAPS:
                feats[FEATURE_TAPS[i]] = x
        x = self.last(x, act="hardswish")
        a = self.aspp1(x, act="relu")
        g = self.aspp2(X.global_avgpool(x), act="sigmoid")
        feats["f4"] = X.channel_scale(a, g)
        return feats

    def forward(self, src_u8: torch.Tensor, state: List[torch.Tensor], *, mode: Optional[int] = None,
                want_fgr_pha: bool = True, out_u8: Optional[torch.Tensor] = None):
        """src_u8 [N,H,W,3] uint8; state r1..r4 (new_state) updated in place -> fgr, pha, composited u8 frame."""
        n, H, W, _ = src_u8.shape
        r = self.cfg.downsample_ratio
        h, w = self.working_size(H, W)
        step = 1.0 if r == 1 else 1.0 / r
        s0 = M.resize_bilinear_ex(src_u8, h, w, step, step, dtype=self.dtype)
        ft = self._encoder(s0)
        s1 = M.avgpool2x_ceil(s0)
        s2 = M.avgpool2x_ceil(s1)
        s3 = M.avgpool2x_ceil(s2)
        r1, r2, r3, r4 = state
        x = ft["f4"]
        self.gru4(x, r4)
        for (conv, gru), skip, s, hstate in zip(self.up, (ft["f3"], ft["f2"], ft["f1"]), (s3, s2, s1), (r3, r2, r1)):
            up = M.resize_bilinear_ex(x, s.shape[1], s.shape[2], 0.5, 0.5)
            x = conv([up, skip, s], act="relu")
            gru(x, hstate)
        up = M.resize_bilinear_ex(x, h, w, 0.5, 0.5)
        hid = self.out1([self.out0([up, s0], act="relu")], act="relu")
        proj = self.project([hid])
        if r != 1:
            stats, mx, my = M.guided_stats(s0, proj)
            A = self.ref3([self.ref2([self.ref1([stats, hid], act="relu")], act="relu")])
            return M.guided_apply(src_u8, A, mx, my, want_fgr_pha=want_fgr_pha, mode=mode, out_u8=out_u8)
        return M.guided_apply(src_u8, proj, None, None, want_fgr_pha=want_fgr_pha, mode=mode, out_u8=out_u8)
