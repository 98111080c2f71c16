"""Drop-in for the reference's customDatasets/perturbations.py: the eight uint8-space Dataset
wrappers `Cls(base_dataset, param)` with `__getitem__(i) -> (fp32 [C,H,W], mask)`.

Random numbers are drawn on the host exactly where the reference draws them (torch.normal on the
image's device; Python `random` for occlusion corners, y before x; salt-and-pepper as a sequence of
(y, x, coin) draws), so seeding reproduces the reference's samples; the image arithmetic runs in the
library's perturbation kernel.  Each wrapper also exposes `.spec()` — the engine.Perturbation the
fused evaluation path uses for whole batches."""
from __future__ import annotations

import random

import numpy as np
import torch
from torch.utils.data import Dataset

from .. import runtime
from ..engine import Perturbation


class _U8Perturbation(Dataset):
    def __init__(self, base_dataset, param):
        super().__init__()
        self.base_dataset = base_dataset
        self.param = param

    def __len__(self):
        return len(self.base_dataset)

    def spec(self, image=None) -> Perturbation:
        raise NotImplementedError

    def __getitem__(self, idx):
        image, mask = self.base_dataset[idx]
        spec = self.spec(image)
        dev = image.device if image.is_cuda else torch.device("cuda", torch.cuda.current_device())
        eng = runtime.engine_for(dev)
        out = eng.perturb(image.to(dev, torch.float32).unsqueeze(0).contiguous(), spec)[0]
        return out.to(image.device), mask


class GaussianPixelNoise(_U8Perturbation):
    """reference :8-50"""

    def __init__(self, base_dataset, standard_deviation=9):
        super().__init__(base_dataset, standard_deviation)
        self.std = standard_deviation

    def spec(self, image=None):
        aux = None
        if image is not None:
            aux = torch.normal(mean=0.0, std=float(self.std), size=image.shape, device=image.device,
                               dtype=torch.float).unsqueeze(0)
        return Perturbation.u8_noise(self.std, aux=aux)


class GaussianBlur(_U8Perturbation):
    """reference :53-124"""

    def __init__(self, base_dataset, num_blur_passes=0):
        super().__init__(base_dataset, num_blur_passes)
        self.num_blur_passes = num_blur_passes

    def spec(self, image=None):
        return Perturbation.u8_blur(self.num_blur_passes)


class ContrastIncrease(_U8Perturbation):
    """reference :126-156"""

    def __init__(self, base_dataset, scale_factor=1.0):
        super().__init__(base_dataset, scale_factor)
        self.scale_factor = scale_factor

    def spec(self, image=None):
        return Perturbation.u8_contrast(self.scale_factor)


class ContrastDecrease(ContrastIncrease):
    """reference :159-198 (same arithmetic as ContrastIncrease)"""


class BrightnessIncrease(_U8Perturbation):
    """reference :200-235"""

    def __init__(self, base_dataset, offset=0):
        super().__init__(base_dataset, offset)
        self.offset = offset

    def spec(self, image=None):
        return Perturbation.u8_brightness(self.offset, increase=True)


class BrightnessDecrease(_U8Perturbation):
    """reference :240-275"""

    def __init__(self, base_dataset, offset=0):
        super().__init__(base_dataset, offset)
        self.offset = offset

    def spec(self, image=None):
        return Perturbation.u8_brightness(self.offset, increase=False)


class OcclusionIncrease(_U8Perturbation):
    """reference :279-327"""

    def __init__(self, base_dataset, square_size=0):
        super().__init__(base_dataset, square_size)
        self.square_size = square_size

    def draw_corner(self, h, w):
        s = self.square_size
        if s > 0 and (h - s) > 0 and (w - s) > 0:
            y0 = random.randint(0, h - s)
            x0 = random.randint(0, w - s)
            return y0, x0
        return -1, -1

    def spec(self, image=None):
        aux = None
        if image is not None:
            aux = torch.tensor([self.draw_corner(image.shape[1], image.shape[2])], dtype=torch.int32)
        return Perturbation.u8_occlusion(self.square_size, aux=aux)


class SaltPepperNoise(_U8Perturbation):
    """reference :330-382"""

    def __init__(self, base_dataset, amount=0.0):
        super().__init__(base_dataset, amount)
        self.amount = amount

    def draw_code(self, h, w):
        code = np.zeros((h, w), dtype=np.uint8)
        if self.amount > 0.0:
            for _ in range(int(round(self.amount * h * w))):
                y = random.randint(0, h - 1)
                x = random.randint(0, w - 1)
                code[y, x] = 1 if random.random() < 0.5 else 2
        return code

    def spec(self, image=None):
        aux = None
        hw = 256 * 256
        if image is not None:
            hw = image.shape[1] * image.shape[2]
            aux = torch.from_numpy(self.draw_code(image.shape[1], image.shape[2])).unsqueeze(0)
        return Perturbation.u8_salt_pepper(self.amount, aux=aux, hw=hw)
