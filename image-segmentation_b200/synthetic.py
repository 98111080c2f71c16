"""Synthetic stand-in for the Oxford-IIIT-Pet test split the reference evaluates on (no network here).

Honours the output contract of customDatasets/datasets.py:137-160: image = decoded uint8 [H,W,3]
(the Dataset divides by 255 -> fp32 CHW in [0,1], values exactly k/255), mask int64/uint8 [H,W] in
{0 background, 1 cat, 2 dog}.  Images are a smooth colour field plus pixel noise; the mask is a
function of the smooth field, so a network can actually learn it (used to give parity tests weights
with real decision margins).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def make_dataset(n: int, h: int = 256, w: int = 256, seed: int = 99):
    """-> (images uint8 [n,h,w,3], masks uint8 [n,h,w]) on CPU, deterministic in (n,h,w,seed)."""
    g = torch.Generator().manual_seed(seed)
    coarse = torch.rand((n, 3, 8, 8), generator=g)
    field = F.interpolate(coarse, size=(h, w), mode="bicubic", align_corners=False)
    noise = torch.randn((n, 3, h, w), generator=g) * 0.08
    img = (field + noise).clamp(0, 1)
    images = (img * 255.0).round().clamp(0, 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous()
    masks = field.argmax(dim=1).to(torch.uint8).contiguous()
    return images, masks


def to_float_nchw(images_u8_nhwc: torch.Tensor) -> torch.Tensor:
    """What CustomImageDataset.__getitem__ hands over (datasets.py:115,157): fp32 CHW = u8/255."""
    return images_u8_nhwc.permute(0, 3, 1, 2).float() / 255.0


def prompt_heatmaps(n: int, h: int = 256, w: int = 256, seed: int = 3, sigma: float = 10.0) -> torch.Tensor:
    """Gaussian prompt heat-maps [n,1,h,w] as customDatasets/datasets.py:599-610 builds them (exp(-d^2 / 2 sigma^2) around
    a seeded click position)."""
    g = torch.Generator().manual_seed(seed)
    cy = torch.randint(0, h, (n,), generator=g).float()
    cx = torch.randint(0, w, (n,), generator=g).float()
    yv, xv = torch.meshgrid(torch.arange(h, dtype=torch.float32), torch.arange(w, dtype=torch.float32), indexing="ij")
    d2 = (xv[None] - cx[:, None, None]) ** 2 + (yv[None] - cy[:, None, None]) ** 2
    return torch.exp(-d2 / (2.0 * sigma ** 2)).unsqueeze(1)


def amplify_bottleneck_path(state_dict, up: float = 16.0, prompt_fusion: float = 32.0, prompt_stem: float = 100.0):
    """With default initialisation the decoder's up-sampled half carries ~1e-6 of a logit, so a test on logits would
    not notice a wrong bottleneck (CLIP fusion, prompt encoder).  Scales `dec*.up.weight`, the prompt half of
    `prompt_fusion.weight` and the prompt encoder's first conv so that those branches move the logits by O(0.1)."""
    sd = dict(state_dict)
    for k in list(sd):
        if k.endswith(".up.weight") and k.startswith("dec"):
            sd[k] = sd[k] * up
    if "prompt_fusion.weight" in sd:
        w = sd["prompt_fusion.weight"].clone()
        w[:, w.shape[1] // 2:] *= prompt_fusion
        sd["prompt_fusion.weight"] = w
        k0 = "prompt_encoder.enc1.block.0.conv.0.weight"
        sd[k0] = sd[k0] * prompt_stem
    return sd


def randomize_batchnorm(state_dict, seed: int = 7):
    """Give every BatchNorm non-trivial affine + running statistics so BN folding is exercised."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k, v in state_dict.items():
        if k.endswith("running_var"):
            out[k] = torch.rand(v.shape, generator=g) + 0.5
        elif k.endswith("running_mean"):
            out[k] = torch.randn(v.shape, generator=g) * 0.1
        elif ".conv.1." in k or ".conv.4." in k:
            if k.endswith("weight"):
                out[k] = torch.rand(v.shape, generator=g) + 0.5
            elif k.endswith("bias"):
                out[k] = torch.randn(v.shape, generator=g) * 0.1
            else:
                out[k] = v.clone()
        else:
            out[k] = v.clone()
    return out


class SyntheticSegmentationDataset(torch.utils.data.Dataset):
    """`Dataset` with the reference's item contract: (fp32 [3,H,W] in [0,1], int64 [H,W])."""

    def __init__(self, n: int = 64, h: int = 256, w: int = 256, seed: int = 99, **_ignored):
        self.images_u8, self.masks_u8 = make_dataset(n, h, w, seed)

    def __len__(self):
        return self.images_u8.shape[0]

    def __getitem__(self, i):
        return self.images_u8[i].permute(2, 0, 1).float() / 255.0, self.masks_u8[i].long()


def make_raw_rows(n: int, h: int = 256, w: int = 256, seed: int = 99):
    """Rows with the schema the reference reads from Oxford-IIIT-Pet-Augmented (customDatasets/datasets.py:107-110):
    {'image': H*W*3 bytes (uint8 HWC), 'mask': H*W bytes (trimap: 38 = cat body, 75 = dog body, 255 = uncertain
    border, 0 = background)}.  Even rows are cats, odd rows dogs; row 2 (if present) has no animal at all."""
    images, cls = make_dataset(n, h, w, seed)
    body = cls != 0
    pad = F.max_pool2d(body.float().unsqueeze(1), 5, stride=1, padding=2).squeeze(1) > 0
    core = -F.max_pool2d(-body.float().unsqueeze(1), 5, stride=1, padding=2).squeeze(1) > 0
    rows = []
    for i in range(n):
        m = torch.zeros((h, w), dtype=torch.uint8)
        m[pad[i]] = 255
        m[core[i]] = 38 if i % 2 == 0 else 75
        if i == 2:
            m[m == 38] = 0
        rows.append({"image": images[i].numpy().tobytes(), "mask": m.numpy().tobytes()})
    return rows
