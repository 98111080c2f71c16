"""ORACLE (test infrastructure, not product): CPU restatement of the reference's evaluation metrics.

  IoU            models/losses.py:38-63     softmax -> argmax -> per-class I,U over the batch (fp32)
  PixelAccuracy  models/losses.py:129-154   3 classes hard-coded, mean over classes present
  Dice           models/losses.py:92-100    1 - smp.losses.DiceLoss("multiclass")(softmax(z), t)
  wrapper dice   models/model_wrappers.py:458   2*IoU/(1+IoU)

Dice depends on segmentation_models_pytorch==0.4.0 (requirements.txt:56), which is NOT under
/root/reference and not installed: `dice_smp` restates its published DiceLoss algorithm
(from_logits=True, smooth=0, eps=1e-7, dims=(0,2), loss masked by class presence, mean over
classes).  No reference test or vector covers it -> Dice is PARITY UNPINNED.  IoU / PixelAccuracy are
pinned against the reference's own modules (tests/test_oracle_pinned.py, tests/golden/).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def confusion_matrix(pred: np.ndarray, target: np.ndarray, n_classes: int = 3) -> np.ndarray:
    """int64 [C,C], rows = target class, cols = predicted class."""
    pred = np.asarray(pred).astype(np.int64).ravel()
    target = np.asarray(target).astype(np.int64).ravel()
    idx = target * n_classes + pred
    return np.bincount(idx, minlength=n_classes * n_classes).reshape(n_classes, n_classes).astype(np.int64)


def iou_from_cm(cm: np.ndarray, eps: float = 1e-6) -> np.float32:
    """losses.py:52-63 from counts: all operands are fp32 tensors in the reference."""
    cm = np.asarray(cm, dtype=np.int64)
    vals = []
    for c in range(cm.shape[0]):
        inter = np.float32(cm[c, c])
        pred_c = np.float32(cm[:, c].sum())
        targ_c = np.float32(cm[c, :].sum())
        union = np.float32(np.float32(pred_c + targ_c) - inter)
        vals.append(np.float32(np.float32(inter + np.float32(eps)) / np.float32(union + np.float32(eps))))
    return torch.mean(torch.tensor(vals, dtype=torch.float32)).numpy()[()]


def pixel_accuracy_from_cm(cm: np.ndarray) -> np.float32:
    """losses.py:141-154: classes 0..2 only, skipped when absent from the targets."""
    cm = np.asarray(cm, dtype=np.int64)
    vals = []
    for c in range(3):
        total = cm[c, :].sum()
        if total > 0:
            vals.append(np.float32(np.float32(cm[c, c]) / np.float32(total)))
    return torch.mean(torch.tensor(vals, dtype=torch.float32)).numpy()[()]


def wrapper_dice_from_iou(iou: np.float32) -> np.float32:
    """model_wrappers.py:458: 2*iou/(1+iou) on 0-dim fp32 tensors."""
    i = torch.tensor(iou, dtype=torch.float32)
    return (2 * i / (1 + i)).numpy()[()]


@torch.no_grad()
def iou_direct(logits: torch.Tensor, targets: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """Straight restatement of losses.py:43-63 (used to pin iou_from_cm)."""
    pred = torch.argmax(F.softmax(logits, dim=1), dim=1)
    out = []
    for c in range(logits.shape[1]):
        p = (pred == c).float()
        t = (targets == c).float()
        inter = (p * t).sum()
        union = p.sum() + t.sum() - inter
        out.append((inter + eps) / (union + eps))
    return torch.mean(torch.stack(out))


@torch.no_grad()
def pixel_accuracy_direct(logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    pred = torch.argmax(F.softmax(logits, dim=1), dim=1)
    out = []
    for c in range(3):
        m = targets == c
        if m.sum() > 0:
            out.append(((pred == targets) & m).float().sum() / m.float().sum())
    return torch.stack(out).mean()


@torch.no_grad()
def dice_sums(logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    """Per class [C,3] fp64: (sum p_c*1[t=c], sum p_c, sum 1[t=c]) with p = softmax(softmax(z)).

    These are the partial sums the CUDA head kernel accumulates per reference batch."""
    p = F.softmax(logits.float(), dim=1)           # losses.py:100
    p = p.log_softmax(dim=1).exp()                  # smp DiceLoss from_logits=True
    c = logits.shape[1]
    onehot = F.one_hot(targets.long(), c).permute(0, 3, 1, 2).to(p.dtype)
    inter = (p * onehot).sum(dim=(0, 2, 3)).double()
    psum = p.sum(dim=(0, 2, 3)).double()
    tsum = onehot.sum(dim=(0, 2, 3)).double()
    return torch.stack([inter, psum, tsum], dim=1)


def dice_from_sums(sums) -> float:
    """smp 0.4.0 DiceLoss(mode='multiclass'): score=(2I)/clamp_min(P+T,1e-7); loss=(1-score)*[T>0];
    mean over classes; reference Dice = 1 - loss (losses.py:100)."""
    s = np.asarray(sums, dtype=np.float64)
    inter, psum, tsum = s[:, 0], s[:, 1], s[:, 2]
    score = 2.0 * inter / np.maximum(psum + tsum, 1e-7)
    loss = (1.0 - score) * (tsum > 0)
    return float(1.0 - loss.mean())


def dice_smp(logits: torch.Tensor, targets: torch.Tensor) -> float:
    return dice_from_sums(dice_sums(logits, targets).numpy())
