"""Drop-in for the evaluation metrics of the reference's models/losses.py: IoU (:38-63),
PixelAccuracy (:129-154), Dice (:92-100).  `m(preds[B,C,H,W] logits, targets[B,H,W]) -> 0-dim tensor`.

The counting (softmax -> argmax -> confusion counts, and the soft-Dice sums) runs in the library's
metrics kernel (ub_metrics); the final ratios are the reference's fp32 formulas (../metrics.py)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import metrics, runtime


def _counts(preds, targets, want_dice=False):
    if not preds.is_cuda:
        raise RuntimeError("unet-b200 metrics run on CUDA only (no CPU fallback)")
    eng = runtime.engine_for(preds.device)
    cm, dice = eng.metrics(preds, targets, ref_batch=preds.shape[0], want_dice=want_dice)
    # the ratios are evaluated on the host in the reference's fp32 arithmetic (bit-reproducible); the
    # reference syncs here too (.item() per batch, model_wrappers.py:461-463)
    return cm[0].cpu(), (dice[0].cpu() if dice is not None else None)


class IoU(nn.Module):
    def __init__(self, eps=1e-6):
        super().__init__()
        self.eps = eps

    def forward(self, preds, targets):
        cm, _ = _counts(preds, targets)
        return metrics.iou_from_cm(cm, self.eps).to(preds.device)


class PixelAccuracy(nn.Module):
    def forward(self, preds, targets):
        cm, _ = _counts(preds, targets)
        return metrics.pixel_accuracy_from_cm(cm).to(preds.device)


class Dice(nn.Module):
    def __init__(self, eps=1e-6):
        super().__init__()

    def forward(self, preds, targets):
        _, sums = _counts(preds, targets, want_dice=True)
        return metrics.dice_from_sums(sums).to(torch.float32).to(preds.device)
