"""UNet model family.

`variant="ref"` is the network the reference actually runs (model/unet_model.py:5-12, model/unet_parts.py:6-76):
conv_block = Conv3x3(bias) -> ReLU -> Conv3x3(bias) -> ReLU, widths 32-64-128-256 (mid 512), one shared 2x2
max-pool, ConvTranspose2d(2,2) up-sampling, skip = center-cropped encoder activation concatenated FIRST,
Conv2d(32,1,1) + sigmoid head.  Its state_dict keys and tensor shapes are identical to the reference's, so
checkpoints are interchangeable (7,760,097 parameters, model/modelsummary.txt:63).

`variant="milesial"` is the upstream model the reference documents in model/modelsummary.txt:153-247 and the
north star names (DoubleConv / Down / Up / OutConv): Conv3x3(bias=False) -> BatchNorm -> ReLU twice, widths
64-128-256-512-1024 (halved at the bottom for `bilinear=True`), ConvTranspose2d or bilinear x2 up-sampling,
Conv2d(64, n_classes, 1) head that returns logits (31,037,698 parameters for n_classes=2).

The modules below are the parameter containers and the torch-op execution path (CPU plumbing, oracle for the
kernels, eval fallback).  On a B200 the training step runs through distributedpytorch_b200.engine, which
reads and writes these same parameters.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


def center_crop(t: torch.Tensor, h: int, w: int) -> torch.Tensor:
    """torchvision.transforms.CenterCrop semantics (model/unet_parts.py:58) on the last two dims; a view when
    nothing is cropped."""
    H, W = t.shape[-2:]
    if H == h and W == w:
        return t
    top = int(round((H - h) / 2.0))
    left = int(round((W - w) / 2.0))
    return t[..., top:top + h, left:left + w]


class conv_block(nn.Module):  # noqa: N801  (name kept: it is part of the reference's state_dict keys)
    def __init__(self, insize: int, outsize: int):
        super().__init__()
        self.conv_block = nn.Sequential(
            nn.Conv2d(insize, outsize, kernel_size=3, padding=1),
            nn.ReLU(),
            nn.Conv2d(outsize, outsize, kernel_size=3, padding=1),
            nn.ReLU(),
        )

    def forward(self, x):
        return self.conv_block(x)


class Encoder(nn.Module):
    def __init__(self, in_channels: int = 3, widths: Sequence[int] = (32, 64, 128, 256)):
        super().__init__()
        chans = [in_channels] + list(widths)
        for i in range(4):
            setattr(self, f"conv{i + 1}", conv_block(chans[i], chans[i + 1]))
        self.maxpool = nn.MaxPool2d(2, 2)

    def forward(self, x) -> Tuple[torch.Tensor, ...]:
        skips = []
        for i in range(4):
            x = getattr(self, f"conv{i + 1}")(x)
            skips.append(x)
            x = self.maxpool(x)
        return (x, skips[3], skips[2], skips[1], skips[0])


class Decoder(nn.Module):
    def __init__(self, widths: Sequence[int] = (32, 64, 128, 256)):
        super().__init__()
        w = list(widths)[::-1]  # 256, 128, 64, 32
        for i in range(4):
            setattr(self, f"conv{i + 1}", conv_block(2 * w[i], w[i]))
        for i in range(4):
            setattr(self, f"deconv{i + 1}", nn.ConvTranspose2d(2 * w[i], w[i], 2, 2))

    def forward(self, x, conv1, conv2, conv3, conv4):
        for i, skip in enumerate((conv1, conv2, conv3, conv4)):
            x = getattr(self, f"deconv{i + 1}")(x)
            skip = center_crop(skip, x.shape[2], x.shape[3])
            x = getattr(self, f"conv{i + 1}")(torch.cat((skip, x), dim=1))
        return x


# --------------------------------------------------------------------------------------------------------
# milesial parts
# --------------------------------------------------------------------------------------------------------
class DoubleConv(nn.Module):
    def __init__(self, in_channels: int, out_channels: int, mid_channels: Optional[int] = None):
        super().__init__()
        mid = mid_channels or out_channels
        self.double_conv = nn.Sequential(
            nn.Conv2d(in_channels, mid, kernel_size=3, padding=1, bias=False),
            nn.BatchNorm2d(mid),
            nn.ReLU(inplace=True),
            nn.Conv2d(mid, out_channels, kernel_size=3, padding=1, bias=False),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
        )

    def forward(self, x):
        return self.double_conv(x)


class Down(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.maxpool_conv = nn.Sequential(nn.MaxPool2d(2), DoubleConv(in_channels, out_channels))

    def forward(self, x):
        return self.maxpool_conv(x)


class Up(nn.Module):
    def __init__(self, in_channels: int, out_channels: int, bilinear: bool = False):
        super().__init__()
        if bilinear:
            self.up = nn.Upsample(scale_factor=2, mode="bilinear", align_corners=True)
            self.conv = DoubleConv(in_channels, out_channels, in_channels // 2)
        else:
            self.up = nn.ConvTranspose2d(in_channels, in_channels // 2, kernel_size=2, stride=2)
            self.conv = DoubleConv(in_channels, out_channels)

    def forward(self, x1, x2):
        x1 = self.up(x1)
        dy, dx = x2.size(2) - x1.size(2), x2.size(3) - x1.size(3)
        if dy or dx:
            x1 = F.pad(x1, [dx // 2, dx - dx // 2, dy // 2, dy - dy // 2])
        return self.conv(torch.cat([x2, x1], dim=1))


class OutConv(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.conv = nn.Conv2d(in_channels, out_channels, kernel_size=1)

    def forward(self, x):
        return self.conv(x)


class UNet(nn.Module):
    """UNet(pipe=False) as in the reference; `variant`, `n_channels`, `n_classes`, `bilinear` select the family
    member.  `pipe=True` only records the request: the pipeline engine (parallel/pipe.py) partitions the model
    into 2..N stages itself instead of hard-wiring cuda:0 / cuda:1 inside forward (model/unet_model.py:14-53)."""

    def __init__(self, pipe: bool = False, variant: str = "ref", n_channels: int = 3, n_classes: Optional[int] = None,
                 bilinear: bool = False):
        super().__init__()
        if variant not in ("ref", "milesial"):
            raise ValueError(f"unknown UNet variant {variant!r}")
        self.variant = variant
        self.n_channels = n_channels
        self.bilinear = bilinear
        self.pipeline = pipe
        if variant == "ref":
            self.n_classes = 1 if n_classes is None else n_classes
            if self.n_classes != 1 or bilinear:
                raise ValueError("variant='ref' is the reference's 1-class ConvTranspose model")
            self.encoder = Encoder(n_channels)
            self.mid = conv_block(256, 512)
            self.decoder = Decoder()
            self.segmap = nn.Conv2d(32, 1, 1)
            self.sigmoid = nn.Sigmoid()
        else:
            self.n_classes = 2 if n_classes is None else n_classes
            factor = 2 if bilinear else 1
            self.inc = DoubleConv(n_channels, 64)
            self.down1 = Down(64, 128)
            self.down2 = Down(128, 256)
            self.down3 = Down(256, 512)
            self.down4 = Down(512, 1024 // factor)
            self.up1 = Up(1024, 512 // factor, bilinear)
            self.up2 = Up(512, 256 // factor, bilinear)
            self.up3 = Up(256, 128 // factor, bilinear)
            self.up4 = Up(128, 64, bilinear)
            self.outc = OutConv(64, self.n_classes)

    @property
    def returns_logits(self) -> bool:
        return self.variant == "milesial"

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.variant == "ref":
            x, conv1, conv2, conv3, conv4 = self.encoder(x)
            x = self.mid(x)
            x = self.decoder(x, conv1, conv2, conv3, conv4)
            return self.sigmoid(self.segmap(x))
        x1 = self.inc(x)
        x2 = self.down1(x1)
        x3 = self.down2(x2)
        x4 = self.down3(x3)
        x5 = self.down4(x4)
        x = self.up1(x5, x4)
        x = self.up2(x, x3)
        x = self.up3(x, x2)
        x = self.up4(x, x1)
        return self.outc(x)

    # ---- stage decomposition used by the pipeline engine (torch-op path) and the cost model -------------
    def stage_functions(self) -> List[Tuple[str, float]]:
        """(block name, forward GFLOP per 640x960 image) in execution order; the pipeline partitioner cuts this
        list into contiguous stages (the reference's fixed split is 35% / 65%)."""
        if self.variant == "ref":
            return [("enc1", 12.39), ("enc2", 16.99), ("enc3", 16.99), ("enc4", 16.99), ("mid", 16.99),
                    ("dec1", 36.50), ("dec2", 36.50), ("dec3", 36.50), ("dec4", 36.50 + 0.04)]
        scale = 4.0
        return [("inc", 12.39 * scale), ("down1", 16.99 * scale), ("down2", 16.99 * scale), ("down3", 16.99 * scale),
                ("down4", 16.99 * scale), ("up1", 36.5 * scale), ("up2", 36.5 * scale), ("up3", 36.5 * scale),
                ("up4", 36.5 * scale)]


    # measured forward + backward milliseconds of each block on one B200 (batch 4, 640x960, this repo's kernels;
    # profiles/step_ops_b4_r1_final.txt and profiles/step_ops_milesial_b4_r2.txt, regenerate with scripts/profile_step.py).  The
    # full-resolution blocks are bandwidth-bound, so time - not FLOPs - is what the pipeline partitioner must balance:
    # by FLOPs dec4 is 16 % of the reference model, by time it is 22 %.
    MEASURED_MS = {
        "ref": [0.616, 0.423, 0.353, 0.334, 0.353, 0.612, 0.572, 0.657, 1.107],
        "milesial": [1.729, 1.341, 1.094, 1.027, 1.248, 2.004, 1.957, 2.107, 3.136],
    }

    def stage_costs(self) -> List[Tuple[str, float]]:
        """(block name, measured fwd+bwd ms) - the cost model of the pipeline partitioner (parallel/pipe.py)."""
        names = [n for n, _ in self.stage_functions()]
        return list(zip(names, self.MEASURED_MS[self.variant]))


def count_parameters(model: nn.Module) -> int:
    return sum(p.numel() for p in model.parameters() if p.requires_grad)
