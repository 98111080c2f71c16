"""ORACLE (test infrastructure, not product): CPU restatement of the reference's perturbations.

(1) uint8-space Dataset wrappers, customDatasets/perturbations.py — restated as exact INTEGER maps in
    numpy (every wrapper quantises to uint8 first, lines 34,89,144,186,226,266,305,358, and returns
    k/255.0f):
      GaussianPixelNoise :31-50   u' = rint(clamp(u + noise, 0, 255))      noise tensor supplied
      GaussianBlur       :85-124  n x { 3x3 zero-padded mean, rint, clamp } == floor((S+4)/9)
      ContrastIncrease/Decrease :140-156,:182-198   LUT[u] = clamp(rint(fl32(u*fl32(s))),0,255)
      BrightnessIncrease :222-235 (clamp then round), BrightnessDecrease :262-275 (round then clamp)
      OcclusionIncrease  :301-327 zero square, corner (y0,x0) supplied (Python `random`, y first)
      SaltPepperNoise    :354-382 per-pixel code map supplied (0 keep, 1 pepper=0, 2 salt=255)
(2) float-space nn.Modules, models/processing_blocks.py:454-592 — restated in numpy fp32.
    RepeatedBlur (:493-496) calls kornia==0.8.0 `filters.box_blur(img,(3,3))` (requirements.txt:20),
    absent from /root/reference and the image: restated from its published algorithm (normalised
    3x3 mean, border_type='reflect', fp32, no re-quantisation) -> RepeatedBlur is PARITY UNPINNED.

Pinned (everything except RepeatedBlur): tests/test_oracle_pinned.py runs the reference's own classes
here and tests/golden/perturb_golden.npz holds their outputs (made by oracle/make_golden.py).
The random-draw replay helpers mirror the reference's RNG consumption order exactly.
"""
from __future__ import annotations

import random

import numpy as np

F32 = np.float32


def quantise(img: np.ndarray) -> np.ndarray:
    """(image*255).round().clamp(0,255).to(uint8); torch.round is half-to-even like np.rint."""
    return np.clip(np.rint(img.astype(F32) * F32(255.0)), 0, 255).astype(np.uint8)


def dequantise(u8: np.ndarray) -> np.ndarray:
    return u8.astype(F32) / F32(255.0)


# ---------------------------------------------------------------- uint8-space (Dataset wrappers)
def u8_gaussian_noise(u8, noise):
    """noise: fp32 array, same shape, ALREADY scaled by std (torch.normal(0,std))."""
    v = u8.astype(F32) + noise.astype(F32)
    return np.rint(np.clip(v, F32(0), F32(255))).astype(np.uint8)


def u8_blur(u8, passes: int):
    """u8: [..., H, W] (channels leading).  Zero padding at the image border on every pass."""
    cur = u8.astype(np.int32)
    for _ in range(int(passes)):
        p = np.pad(cur, [(0, 0)] * (cur.ndim - 2) + [(1, 1), (1, 1)])
        h, w = cur.shape[-2:]
        s = np.zeros_like(cur)
        for dy in range(3):
            for dx in range(3):
                s += p[..., dy:dy + h, dx:dx + w]
        cur = (s + 4) // 9
    return cur.astype(np.uint8)


def contrast_lut(scale: float) -> np.ndarray:
    u = np.arange(256, dtype=F32)
    return np.clip(np.rint(u * F32(scale)), 0, 255).astype(np.uint8)


def u8_contrast(u8, scale):
    return contrast_lut(scale)[u8]


def brightness_lut(offset: float, increase: bool) -> np.ndarray:
    u = np.arange(256, dtype=F32)
    if increase:
        return np.rint(np.clip(u + F32(float(offset)), 0, 255)).astype(np.uint8)
    return np.clip(np.rint(u - F32(float(offset))), 0, 255).astype(np.uint8)


def u8_brightness(u8, offset, increase=True):
    return brightness_lut(offset, increase)[u8]


def u8_occlusion(u8, size: int, y0: int, x0: int):
    """u8: [C,H,W]; (y0,x0) < 0 means 'no square drawn'."""
    out = u8.copy()
    if size > 0 and y0 >= 0 and x0 >= 0:
        out[:, y0:y0 + size, x0:x0 + size] = 0
    return out


def u8_salt_pepper(u8, code):
    """code: uint8 [H,W]: 0 keep, 1 -> 0 (pepper), 2 -> 255 (salt); all channels."""
    out = u8.copy()
    out[:, code == 1] = 0
    out[:, code == 2] = 255
    return out


def replay_occlusion_corner(size: int, h: int, w: int, rng=random):
    """perturbations.py:309-319: draws happen only if size>0 and both maxima > 0; y0 before x0."""
    if size > 0 and (h - size) > 0 and (w - size) > 0:
        y0 = rng.randint(0, h - size)
        x0 = rng.randint(0, w - size)
        return y0, x0
    return -1, -1


def replay_salt_pepper_code(amount: float, h: int, w: int, rng=random) -> np.ndarray:
    """perturbations.py:362-377: round(amount*H*W) sequential draws (y, x, coin); later draws win."""
    code = np.zeros((h, w), dtype=np.uint8)
    if amount > 0.0:
        for _ in range(int(round(amount * h * w))):
            y = rng.randint(0, h - 1)
            x = rng.randint(0, w - 1)
            code[y, x] = 1 if rng.random() < 0.5 else 2
    return code


# ---------------------------------------------------------------- float-space (nn.Modules)
def f_gaussian_noise(img, randn, std):
    """processing_blocks.py:471-474; randn = torch.randn_like(img) supplied."""
    scale = F32(std / 255.0)
    return np.clip(img.astype(F32) + randn.astype(F32) * scale, F32(0), F32(1))


def f_box_blur(img, times: int):
    """kornia box_blur((3,3)), reflect border, fp32; img [...,H,W].  PARITY UNPINNED."""
    cur = img.astype(F32)
    ninth = F32(1.0) / F32(9.0)
    for _ in range(int(times)):
        p = np.pad(cur, [(0, 0)] * (cur.ndim - 2) + [(1, 1), (1, 1)], mode="reflect")
        h, w = cur.shape[-2:]
        acc = np.zeros_like(cur)
        for dy in range(3):
            for dx in range(3):
                acc = acc + p[..., dy:dy + h, dx:dx + w] * ninth
        cur = acc.astype(F32)
    return cur


def f_contrast(img, factor):
    return np.clip(img.astype(F32) * F32(factor), F32(0), F32(1))


def f_brightness(img, offset):
    """:535 divides in Python double, :538 adds as an fp32 scalar."""
    return np.clip(img.astype(F32) + F32(offset / 255.0), F32(0), F32(1))


def f_occlusion(img, size, corners):
    """img [B,C,H,W]; corners [B,2] as (y,x).  The reference mutates its input in place (:557-563)."""
    out = img.copy()
    for i, (y, x) in enumerate(corners):
        out[i, :, y:y + size, x:x + size] = 0.0
    return out


def replay_f_occlusion_corners(b, size, h, w, rng=random):
    """:559-561 draws x first, then y, per image; each only if the dimension exceeds the size."""
    out = []
    for _ in range(b):
        x = rng.randint(0, w - size) if w > size else 0
        y = rng.randint(0, h - size) if h > size else 0
        out.append((y, x))
    return out


def f_salt_pepper(img, rand, amount):
    """:586-592; rand = torch.rand((b,1,h,w)) supplied.  salt -> 1.0, pepper -> 0.0."""
    rand = rand.astype(F32)
    salt = (rand < F32(amount / 2)).astype(F32)
    pepper = (rand > F32(1 - amount / 2)).astype(F32)
    mask = F32(1) - salt - pepper
    return (img.astype(F32) * mask + salt).astype(F32)
