"""Drop-in for the parts of the reference's models/processing_blocks.py that the robustness-evaluation
path uses: the four conv blocks (:21-133) and the six float-space perturbation modules (:454-592).

The conv blocks are PARAMETER CONTAINERS with the reference's sub-module names, shapes and default
initialisation, so checkpoints (`state_dict` keys such as `enc1.block.0.conv.0.weight`) load unchanged.
They do not compute on their own: the owning network (models/UNet.py) runs all of them in one pass of
the sm_100a library.  The perturbation modules call the library's perturbation kernels.
"""
from __future__ import annotations

import random

import torch
import torch.nn as nn

from .. import runtime
from ..engine import Perturbation


def _no_standalone(name):
    raise RuntimeError(
        f"{name} is a parameter container in unet-b200; run it through UNet/LargeUNet.forward "
        "(the fused sm_100a path). There is no per-block PyTorch fallback.")


class ConvBlock(nn.Module):
    """(conv3x3 pad1 -> BatchNorm2d -> ReLU) x 2; reference processing_blocks.py:21-52."""

    def __init__(self, in_channels, out_channels, kernel_size=3, padding=1):
        super().__init__()
        if kernel_size != 3 or padding != 1:
            raise ValueError("unet-b200 implements the reference's 3x3 / padding 1 block only")
        self.conv = nn.Sequential(
            nn.Conv2d(in_channels, out_channels, kernel_size, padding=padding),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
            nn.Conv2d(out_channels, out_channels, kernel_size, padding=padding),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
        )

    def forward(self, x):
        _no_standalone("ConvBlock")


class ConvBlockDownsample(nn.Module):
    """ConvBlock followed by MaxPool2d(2,2); reference processing_blocks.py:54-77."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.block = nn.Sequential(ConvBlock(in_channels, out_channels), nn.MaxPool2d(kernel_size=2, stride=2))

    def forward(self, x):
        _no_standalone("ConvBlockDownsample")


class ConvBlockUpsampleSkip(nn.Module):
    """ConvTranspose2d(k2,s2) -> bilinear resize to the skip -> concat -> ConvBlock; reference :79-109."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.up = nn.ConvTranspose2d(in_channels, out_channels, kernel_size=2, stride=2)
        self.conv = ConvBlock(out_channels * 2, out_channels)

    def forward(self, x, skip):
        _no_standalone("ConvBlockUpsampleSkip")


class ConvBlockUpsample(nn.Module):
    """ConvTranspose2d(k2,s2) -> ConvBlock (no skip); reference :111-133.  Container only."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.up = nn.ConvTranspose2d(in_channels, out_channels, kernel_size=2, stride=2)
        self.conv = ConvBlock(out_channels, out_channels)

    def forward(self, x):
        _no_standalone("ConvBlockUpsample")


# --------------------------------------------------------------------------- perturbation modules
class _DevicePerturbation(nn.Module):
    """Common part: `forward(img[B,3,H,W] fp32 on cuda) -> img`, computed by ub_perturb_images.

    `spec(img)` returns the engine.Perturbation for this call; random families draw their numbers
    exactly where the reference draws them (so seeding `random` / `torch` reproduces its output)."""

    def spec(self, img) -> Perturbation:
        raise NotImplementedError

    def forward(self, img):
        if not img.is_cuda:
            raise RuntimeError(f"{type(self).__name__}: unet-b200 perturbations run on CUDA only")
        eng = runtime.engine_for(img.device)
        return eng.perturb(img.float().contiguous(), self.spec(img))


class GaussianPixelNoise(_DevicePerturbation):
    """clamp(img + randn*std/255, 0, 1); reference :454-474.  The normal draw stays `torch.randn_like`
    (same generator, same order as the reference); the arithmetic runs in the CUDA kernel."""

    def __init__(self, std):
        super().__init__()
        self.std = std

    def spec(self, img):
        return Perturbation.f_noise(self.std, aux=torch.randn_like(img))


class RepeatedBlur(_DevicePerturbation):
    """`times` passes of kornia.filters.box_blur(img,(3,3)) (reflect border); reference :477-496."""

    def __init__(self, times):
        super().__init__()
        self.times = times

    def spec(self, img):
        return Perturbation.f_blur(self.times)


class ContrastChange(_DevicePerturbation):
    """clamp(img*factor, 0, 1); reference :499-518."""

    def __init__(self, factor):
        super().__init__()
        self.factor = factor

    def spec(self, img):
        return Perturbation.f_contrast(self.factor)


class BrightnessChange(_DevicePerturbation):
    """clamp(img + offset/255, 0, 1); reference :521-539."""

    def __init__(self, offset):
        super().__init__()
        self.offset = offset / 255.0

    def spec(self, img):
        p = Perturbation.f_brightness(0.0)
        p.struct.fparam = float(self.offset)
        return p


class Occlusion(_DevicePerturbation):
    """Zero a size x size square per image, corner from Python `random` (x drawn before y);
    reference :542-563.  Like the reference it writes INTO its input tensor and returns it."""

    def __init__(self, size):
        super().__init__()
        self.size = size

    def corners(self, b, h, w):
        out = []
        for _ in range(b):
            x = random.randint(0, w - self.size) if w > self.size else 0
            y = random.randint(0, h - self.size) if h > self.size else 0
            out.append((y, x))
        return torch.tensor(out, dtype=torch.int32).reshape(b, 2)

    def spec(self, img):
        b, _, h, w = img.shape
        return Perturbation.f_occlusion(self.size, aux=self.corners(b, h, w))

    def forward(self, img):
        out = super().forward(img)
        if img.dtype == torch.float32 and img.is_contiguous():
            img.copy_(out)
            return img
        return out


class SaltAndPepper(_DevicePerturbation):
    """One uniform draw per pixel: < amount/2 -> 1.0, > 1-amount/2 -> 0.0; reference :565-592."""

    def __init__(self, amount):
        super().__init__()
        self.amount = amount

    def spec(self, img):
        b, _, h, w = img.shape
        return Perturbation.f_salt_pepper(self.amount, aux=torch.rand((b, 1, h, w), device=img.device))
