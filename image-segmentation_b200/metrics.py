"""Host-side scalar metrics from per-batch confusion matrices / Dice sums, with the reference's own
fp32 formulas so the numbers match it bit for bit given identical counts.

  IoU            models/losses.py:52-63    (I+eps)/(U+eps) per class over the batch, mean over classes
  PixelAccuracy  models/losses.py:141-154  classes 0..2, mean over the classes present in the targets
  wrapper Dice   models/model_wrappers.py:458   2*IoU/(1+IoU)
  soft Dice      models/losses.py:98-100 + segmentation_models_pytorch 0.4.0 DiceLoss('multiclass')
"""
from __future__ import annotations

import torch


def iou_from_cm(cm: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """cm int64 [..., C, C] (rows target, cols prediction) -> fp32 [...]."""
    cm = cm.to(torch.int64)
    inter = torch.diagonal(cm, dim1=-2, dim2=-1).to(torch.float32)
    pred = cm.sum(dim=-2).to(torch.float32)
    targ = cm.sum(dim=-1).to(torch.float32)
    union = pred + targ - inter
    return ((inter + eps) / (union + eps)).mean(dim=-1)


def pixel_accuracy_from_cm(cm: torch.Tensor) -> torch.Tensor:
    cm = cm.to(torch.int64)[..., :3, :]
    correct = torch.diagonal(cm, dim1=-2, dim2=-1)[..., :3].to(torch.float32)
    total = cm.sum(dim=-1).to(torch.float32)
    present = total > 0
    acc = torch.where(present, correct / total.clamp_min(1), torch.zeros_like(total))
    # mean over present classes, accumulated in class order like torch.stack(...).mean()
    return acc.sum(dim=-1) / present.sum(dim=-1).to(torch.float32)


def wrapper_dice(iou: torch.Tensor) -> torch.Tensor:
    return 2 * iou / (1 + iou)


def dice_from_sums(sums: torch.Tensor, eps: float = 1e-7) -> torch.Tensor:
    """sums f64 [..., C, 3] = (sum p*1[t=c], sum p, sum 1[t=c]) -> reference Dice score [...]."""
    inter, psum, tsum = sums[..., 0], sums[..., 1], sums[..., 2]
    score = 2.0 * inter / (psum + tsum).clamp_min(eps)
    loss = (1.0 - score) * (tsum > 0).to(sums.dtype)
    return 1.0 - loss.mean(dim=-1)
