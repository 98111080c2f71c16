"""Fused robustness evaluation: the body of the reference's two hot loops
(models/model_wrappers.py:440-468 and scripts/robustness_evaluation.py:105-127) for a whole
perturbation level at once, sharded by REFERENCE BATCH over the GPUs of one box.

The reference computes every metric per DataLoader batch and averages over batches
(model_wrappers.py:461-468, robustness_evaluation.py:122-124), so the unit of sharding is the batch:
rank r owns a contiguous range of batches, keeps the corresponding slice of the dataset resident on
its GPU (uint8 NHWC images + uint8 masks), fills its rows of the [n_batches,C,C] int64 confusion
tensor, and ONE all-reduce (NCCL, sum) per level combines the rows.  No other collective exists on
this path.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.distributed as dist

from . import metrics
from .engine import Perturbation


def batch_range(n_batches: int, rank: int, world: int):
    """Contiguous, balanced split of reference batches: the first (n_batches % world) ranks get one more."""
    base, extra = divmod(n_batches, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


class RobustnessEvaluator:
    def __init__(self, model, images_u8: torch.Tensor, masks_u8: torch.Tensor, ref_batch: int, device=None,
                 rank: Optional[int] = None, world_size: Optional[int] = None, max_microbatch: Optional[int] = None):
        """images_u8 [N,H,W,3] uint8, masks_u8 [N,H,W] uint8 — the WHOLE evaluation set (host or device);
        only this rank's batches are uploaded."""
        if rank is None:
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        if world_size is None:
            world_size = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank, self.world = rank, world_size
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.model = model
        self.ref_batch = int(ref_batch)
        self.n_total = images_u8.shape[0]
        self.n_batches = math.ceil(self.n_total / self.ref_batch)
        self.b0, self.b1 = batch_range(self.n_batches, rank, world_size)
        i0, i1 = self.b0 * self.ref_batch, min(self.b1 * self.ref_batch, self.n_total)
        self.i0, self.i1 = i0, i1
        self.images = images_u8[i0:i1].to(self.device, non_blocking=True).contiguous()
        self.masks = masks_u8[i0:i1].to(self.device, non_blocking=True).contiguous()
        self.C = model._n_classes
        self.engine = model.engine(self.device, max_microbatch)
        self.cm = torch.zeros((self.n_batches, self.C, self.C), device=self.device, dtype=torch.int64)
        self.dice = torch.zeros((self.n_batches, self.C, 3), device=self.device, dtype=torch.float64)

    @property
    def n_local(self) -> int:
        return self.i1 - self.i0

    def upload(self, images_u8: torch.Tensor, masks_u8: torch.Tensor):
        """Host -> device copy of this rank's slice into the resident buffers (async on the current stream
        when the host tensors are pinned).  Used when the evaluation set changes between levels."""
        self.images.copy_(images_u8[self.i0:self.i1], non_blocking=True)
        self.masks.copy_(masks_u8[self.i0:self.i1], non_blocking=True)

    def upload_async(self, images_u8: torch.Tensor, masks_u8: torch.Tensor):
        """Double-buffered upload: copies this rank's slice of the NEXT evaluation set into the back buffers on a
        side stream while the current level computes.  `swap()` makes them current (and orders the compute stream
        after the copy).  Host tensors should be pinned."""
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream(self.device)
            self._back = (torch.empty_like(self.images), torch.empty_like(self.masks))
            self._copy_done = torch.cuda.Event()
            self._front_free = torch.cuda.Event()
            self._front_free.record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self._copy_stream):
            # the back buffers were the front buffers two swaps ago: their last reader has been recorded
            self._copy_stream.wait_event(self._front_free)
            self._back[0].copy_(images_u8[self.i0:self.i1], non_blocking=True)
            self._back[1].copy_(masks_u8[self.i0:self.i1], non_blocking=True)
            self._copy_done.record(self._copy_stream)

    def swap(self):
        """Make the buffers filled by `upload_async` the current evaluation set."""
        cur = torch.cuda.current_stream(self.device)
        self._front_free.record(cur)           # everything launched so far may still read the old front buffers
        cur.wait_event(self._copy_done)
        (self.images, self.masks), self._back = self._back, (self.images, self.masks)

    def local_aux(self, perturbation: Optional[Perturbation]):
        """Parity mode hands aux for the whole dataset; keep this rank's rows."""
        if perturbation is None or perturbation.aux is None:
            return perturbation
        aux = perturbation.aux
        if aux.shape[0] == self.n_total and self.n_total != self.n_local:
            p = Perturbation()
            p.struct, p.name = perturbation.struct, perturbation.name
            p.aux = aux[self.i0:self.i1]
            return p
        return perturbation

    def _eval(self, images, masks, spec: Optional[Perturbation], want_dice, out_cm, out_dice):
        """One engine call over `images` (a whole number of reference batches, except possibly the last)."""
        if self.engine.arch != "clip":
            self.engine.eval_batches(images, masks, spec, self.ref_batch, want_dice=want_dice, out_cm=out_cm,
                                     out_dice=out_dice)
            return
        # ClipUnet (models/CLIP_models.py:114-131): the CLIP embedding is taken from the image the network sees, i.e.
        # AFTER the perturbation, so the perturbed batch is materialised, embedded, and then evaluated unperturbed.
        step = max(1, self.engine.max_microbatch // self.ref_batch) * self.ref_batch
        n = images.shape[0]
        for i0 in range(0, n, step):
            i1 = min(i0 + step, n)
            sub = Perturbation()
            if spec is not None:
                sub.struct, sub.name = type(spec.struct).from_buffer_copy(spec.struct), spec.name
                sub.aux = spec.aux[i0:i1] if spec.aux is not None else None
                sub.struct.seed = (spec.struct.seed + 0x9E3779B97F4A7C15 * i0) & (2 ** 64 - 1)  # device RNG: new stream
            x = self.engine.perturb(images[i0:i1].contiguous(), sub)
            feats = self.model.clip_feature_extractor(x)
            b0, b1 = i0 // self.ref_batch, math.ceil(i1 / self.ref_batch)
            self.engine.eval_batches(x, masks[i0:i1].contiguous(), None, self.ref_batch, want_dice=want_dice,
                                     out_cm=out_cm[b0:b1], out_dice=out_dice[b0:b1] if out_dice is not None else None,
                                     clip_feats=feats)

    def run_level(self, perturbation: Optional[Perturbation], want_dice: bool = False, reduce: bool = True):
        """-> (cm int64 [n_batches,C,C], dice_sums f64 [n_batches,C,3] or None), complete on every rank."""
        self.cm.zero_()
        if want_dice:
            self.dice.zero_()
        if self.n_local > 0:
            self._eval(self.images, self.masks, self.local_aux(perturbation), want_dice, self.cm[self.b0:self.b1],
                       self.dice[self.b0:self.b1] if want_dice else None)
        if reduce and self.world > 1:
            dist.all_reduce(self.cm, op=dist.ReduceOp.SUM)
            if want_dice:
                dist.all_reduce(self.dice, op=dist.ReduceOp.SUM)
        return self.cm, (self.dice if want_dice else None)

    def run_level_per_batch(self, spec_for_batch, want_dice: bool = False, chunk_batches: int = 16):
        """Like run_level, but the perturbation's random numbers are drawn batch by batch in loader
        order through `spec_for_batch(shape_probe[b,3,H,W]) -> Perturbation` (what the reference's
        nn.Module augmentations do inside the loop, model_wrappers.py:450), `chunk_batches` at a time."""
        self.cm.zero_()
        if want_dice:
            self.dice.zero_()
        h, w = self.images.shape[1:3]
        for cb0 in range(self.b0, self.b1, chunk_batches):
            cb1 = min(cb0 + chunk_batches, self.b1)
            lo = cb0 * self.ref_batch - self.i0
            hi = min(cb1 * self.ref_batch, self.n_total) - self.i0
            specs = []
            for b in range(cb0, cb1):
                nb = min((b + 1) * self.ref_batch, self.n_total) - b * self.ref_batch
                specs.append(spec_for_batch(torch.empty((nb, 3, h, w), device=self.device, dtype=torch.float32)))
            spec = specs[0]
            if spec.aux is not None:
                merged = Perturbation()
                merged.struct, merged.name = spec.struct, spec.name
                merged.aux = torch.cat([s.aux.to(self.device) for s in specs], dim=0)
                spec = merged
            self._eval(self.images[lo:hi], self.masks[lo:hi], spec, want_dice, self.cm[cb0:cb1],
                       self.dice[cb0:cb1] if want_dice else None)
        if self.world > 1:
            dist.all_reduce(self.cm, op=dist.ReduceOp.SUM)
            if want_dice:
                dist.all_reduce(self.dice, op=dist.ReduceOp.SUM)
        return self.cm, (self.dice if want_dice else None)

    # ---- scalars exactly as the two reference drivers report them
    @staticmethod
    def wrapper_scores(cm: torch.Tensor):
        """model_wrappers.py:456-468: per-batch IoU, PixelAccuracy, 2*IoU/(1+IoU); Python-float means."""
        cm = cm.cpu()
        iou = metrics.iou_from_cm(cm)
        pa = metrics.pixel_accuracy_from_cm(cm)
        dc = metrics.wrapper_dice(iou)
        n = cm.shape[0]
        return (sum(float(v) for v in iou) / n, sum(float(v) for v in pa) / n, sum(float(v) for v in dc) / n)

    @staticmethod
    def script_score(dice_sums: torch.Tensor):
        """robustness_evaluation.py:121-124: per-batch soft Dice (fp32 .item()), mean over batches."""
        d = metrics.dice_from_sums(dice_sums.cpu()).to(torch.float32)
        return sum(float(v) for v in d) / d.shape[0]
