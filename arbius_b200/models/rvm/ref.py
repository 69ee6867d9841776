"""Robust Video Matting (``rvm_mobilenetv3``) — eager restatement with the upstream module/key layout.

The reference only names the container (templates/robust_video_matting.json: ``arielreplicate/robust_video_matting``,
git PeterL1n/RobustVideoMatting); SURVEY §2.7 lists the architecture: MobileNetV3-Large encoder at 1/2..1/16,
LR-ASPP (-> 128), a recurrent decoder with a ConvGRU at each of four scales (states r1..r4 carried frame to frame),
a 16->4 projection (foreground residual + alpha) and a deep-guided-filter refiner back to full resolution.
Keys follow ``MattingNetwork``: backbone.features.*, aspp.*, decoder.decode{4,3,2,1,0}.*, project_mat.conv,
refiner.{box_filter,conv}.*
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


@dataclass
class RvmConfig:
    downsample_ratio: float = 0.25          # ~1080p -> 270x480 working resolution
    width_mult: float = 1.0

    @staticmethod
    def tiny() -> "RvmConfig":
        return RvmConfig(downsample_ratio=0.5)


class MobileNetV3LargeEncoder(nn.Module):
    """torchvision mobilenet_v3_large features with the last three stages dilated (output stride 16)."""

    def __init__(self):
        super().__init__()
        from torchvision.models import mobilenet_v3_large
        net = mobilenet_v3_large(weights=None, dilated=True)
        self.features = net.features

    def forward(self, x):
        f = self.features
        mean = torch.tensor(IMAGENET_MEAN, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
        std = torch.tensor(IMAGENET_STD, dtype=x.dtype, device=x.device).view(1, 3, 1, 1)
        x = (x - mean) / std
        x = f[0](x)
        x = f[1](x)
        f1 = x
        x = f[2](x)
        x = f[3](x)
        f2 = x
        x = f[4](x)
        x = f[5](x)
        x = f[6](x)
        f3 = x
        for i in range(7, 17):
            x = f[i](x)
        return f1, f2, f3, x


class LRASPP(nn.Module):
    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.aspp1 = nn.Sequential(nn.Conv2d(cin, cout, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(True))
        self.aspp2 = nn.Sequential(nn.AdaptiveAvgPool2d(1), nn.Conv2d(cin, cout, 1, bias=False), nn.Sigmoid())

    def forward(self, x):
        return self.aspp1(x) * self.aspp2(x)


class ConvGRU(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.channels = c
        self.ih = nn.Sequential(nn.Conv2d(2 * c, 2 * c, 3, padding=1), nn.Sigmoid())
        self.hh = nn.Sequential(nn.Conv2d(2 * c, c, 3, padding=1), nn.Tanh())

    def forward(self, x, h):
        if h is None:
            h = torch.zeros_like(x)
        r, z = self.ih(torch.cat([x, h], dim=1)).split(self.channels, dim=1)
        c = self.hh(torch.cat([x, r * h], dim=1))
        h = (1 - z) * h + z * c
        return h, h


class BottleneckBlock(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.channels = c
        self.gru = ConvGRU(c // 2)

    def forward(self, x, r):
        a, b = x.split(self.channels // 2, dim=1)
        b, r = self.gru(b, r)
        return torch.cat([a, b], dim=1), r


class UpsamplingBlock(nn.Module):
    def __init__(self, cin: int, cskip: int, csrc: int, cout: int):
        super().__init__()
        self.out_channels = cout
        self.upsample = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=False)
        self.conv = nn.Sequential(nn.Conv2d(cin + cskip + csrc, cout, 3, 1, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(True))
        self.gru = ConvGRU(cout // 2)

    def forward(self, x, f, s, r):
        x = self.upsample(x)[:, :, :s.size(2), :s.size(3)]
        x = self.conv(torch.cat([x, f, s], dim=1))
        a, b = x.split(self.out_channels // 2, dim=1)
        b, r = self.gru(b, r)
        return torch.cat([a, b], dim=1), r


class OutputBlock(nn.Module):
    def __init__(self, cin: int, csrc: int, cout: int):
        super().__init__()
        self.upsample = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=False)
        self.conv = nn.Sequential(nn.Conv2d(cin + csrc, cout, 3, 1, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(True),
                                  nn.Conv2d(cout, cout, 3, 1, 1, bias=False), nn.BatchNorm2d(cout), nn.ReLU(True))

    def forward(self, x, s):
        x = self.upsample(x)[:, :, :s.size(2), :s.size(3)]
        return self.conv(torch.cat([x, s], dim=1))


class RecurrentDecoder(nn.Module):
    def __init__(self, feat: Tuple[int, ...] = (16, 24, 40, 128), dec: Tuple[int, ...] = (80, 40, 32, 16)):
        super().__init__()
        self.avgpool = nn.AvgPool2d(2, 2, count_include_pad=False, ceil_mode=True)
        self.decode4 = BottleneckBlock(feat[3])
        self.decode3 = UpsamplingBlock(feat[3], feat[2], 3, dec[0])
        self.decode2 = UpsamplingBlock(dec[0], feat[1], 3, dec[1])
        self.decode1 = UpsamplingBlock(dec[1], feat[0], 3, dec[2])
        self.decode0 = OutputBlock(dec[2], 3, dec[3])

    def forward(self, s0, f1, f2, f3, f4, r1, r2, r3, r4):
        s1 = self.avgpool(s0)
        s2 = self.avgpool(s1)
        s3 = self.avgpool(s2)
        x4, r4 = self.decode4(f4, r4)
        x3, r3 = self.decode3(x4, f3, s3, r3)
        x2, r2 = self.decode2(x3, f2, s2, r2)
        x1, r1 = self.decode1(x2, f1, s1, r1)
        x0 = self.decode0(x1, s0)
        return x0, r1, r2, r3, r4


class Projection(nn.Module):
    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, 1)

    def forward(self, x):
        return self.conv(x)


class DeepGuidedFilterRefiner(nn.Module):
    def __init__(self, hid: int = 16):
        super().__init__()
        self.box_filter = nn.Conv2d(4, 4, 3, padding=1, bias=False, groups=4)
        self.box_filter.weight.data[...] = 1 / 9
        self.conv = nn.Sequential(nn.Conv2d(4 * 2 + hid, hid, 1, bias=False), nn.BatchNorm2d(hid), nn.ReLU(True),
                                  nn.Conv2d(hid, hid, 1, bias=False), nn.BatchNorm2d(hid), nn.ReLU(True),
                                  nn.Conv2d(hid, 4, 1, bias=True))

    def forward(self, fine_src, base_src, base_fgr, base_pha, base_hid):
        fine_x = torch.cat([fine_src, fine_src.mean(1, keepdim=True)], dim=1)
        base_x = torch.cat([base_src, base_src.mean(1, keepdim=True)], dim=1)
        base_y = torch.cat([base_fgr, base_pha], dim=1)
        mean_x = self.box_filter(base_x)
        mean_y = self.box_filter(base_y)
        cov_xy = self.box_filter(base_x * base_y) - mean_x * mean_y
        var_x = self.box_filter(base_x * base_x) - mean_x * mean_x
        A = self.conv(torch.cat([cov_xy, var_x, base_hid], dim=1))
        b = mean_y - A * mean_x
        H, W = fine_src.shape[2:]
        A = F.interpolate(A, (H, W), mode="bilinear", align_corners=False)
        b = F.interpolate(b, (H, W), mode="bilinear", align_corners=False)
        out = A * fine_x + b
        return out.split([3, 1], dim=1)


class MattingNetwork(nn.Module):
    def __init__(self, cfg: RvmConfig = RvmConfig()):
        super().__init__()
        self.cfg = cfg
        self.backbone = MobileNetV3LargeEncoder()
        self.aspp = LRASPP(960, 128)
        self.decoder = RecurrentDecoder()
        self.project_mat = Projection(16, 4)
        self.refiner = DeepGuidedFilterRefiner()

    def forward(self, src, r1=None, r2=None, r3=None, r4=None, downsample_ratio: Optional[float] = None):
        """src [B, 3, H, W] in [0,1] (one frame per call; recurrent state in/out) -> fgr, pha, r1..r4."""
        ratio = self.cfg.downsample_ratio if downsample_ratio is None else downsample_ratio
        if ratio != 1:
            src_sm = F.interpolate(src, scale_factor=ratio, mode="bilinear", align_corners=False, recompute_scale_factor=False)
        else:
            src_sm = src
        f1, f2, f3, f4 = self.backbone(src_sm)
        f4 = self.aspp(f4)
        hid, r1, r2, r3, r4 = self.decoder(src_sm, f1, f2, f3, f4, r1, r2, r3, r4)
        fgr_res, pha = self.project_mat(hid).split([3, 1], dim=1)
        if ratio != 1:
            fgr_res, pha = self.refiner(src, src_sm, fgr_res, pha, hid)
        fgr = (fgr_res + src).clamp(0.0, 1.0)
        pha = pha.clamp(0.0, 1.0)
        return fgr, pha, r1, r2, r3, r4


def init_rvm_(m: MattingNetwork, seed: int = 0) -> MattingNetwork:
    """Random init incl. non-trivial BatchNorm statistics (so BN folding in the native engine is exercised)."""
    g = torch.Generator().manual_seed(seed)
    for name, mod in sorted(m.named_modules()):
        if isinstance(mod, nn.BatchNorm2d):
            with torch.no_grad():
                mod.weight.copy_(1.0 + 0.1 * torch.randn(mod.weight.shape, generator=g))
                mod.bias.copy_(0.05 * torch.randn(mod.bias.shape, generator=g))
                mod.running_mean.copy_(0.05 * torch.randn(mod.running_mean.shape, generator=g))
                mod.running_var.copy_(1.0 + 0.1 * torch.rand(mod.running_var.shape, generator=g))
        elif isinstance(mod, (nn.Conv2d, nn.Linear)) and name != "refiner.box_filter":
            with torch.no_grad():
                fan_in = mod.weight[0].numel()
                mod.weight.copy_(torch.randn(mod.weight.shape, generator=g) * (1.2 / fan_in ** 0.5))
                if mod.bias is not None:
                    mod.bias.copy_(0.02 * torch.randn(mod.bias.shape, generator=g))
    return m.eval().requires_grad_(False)
