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


def batch_range(n_batches: int, rank: int, world: int, weights=None):
    """Contiguous split of reference batches.  Default: balanced, the first (n_batches % world) ranks get one more.
    `weights` (one positive number per rank, e.g. images/s measured during warm-up): shares proportional to them —
    the boards of one box settle at different power-capped clocks, and a level ends when the slowest rank does."""
    if weights is None:
        base, extra = divmod(n_batches, world)
        start = rank * base + min(rank, extra)
        return start, start + base + (1 if rank < extra else 0)
    w = [float(x) for x in weights]
    if len(w) != world or min(w) <= 0:
        raise ValueError("weights: one positive number per rank")
    total = sum(w)
    # largest-remainder rounding of the cumulative shares: deterministic, identical on every rank
    edges = [0]
    acc = 0.0
    for r in range(world):
        acc += w[r]
        edges.append(int(round(n_batches * acc / total)))
    edges[-1] = n_batches
    for r in range(1, world + 1):
        edges[r] = max(edges[r], edges[r - 1])
    return edges[rank], edges[rank + 1]


class PendingLevel:
    """Result of `RobustnessEvaluator.submit_level`: the counts of one level, whose all-reduce may still be in flight
    on NCCL's stream.  `wait()` orders the CURRENT stream after it and returns (cm, dice_sums or None)."""

    def __init__(self, cm, dice, works):
        self.cm, self.dice, self._works = cm, dice, works

    def wait(self):
        for w in self._works:
            w.wait()
        self._works = []
        return self.cm, self.dice


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
        self.C = model._n_classes
        self._set_range(images_u8, masks_u8, None)
        # Precondition of the fused counts (include/unet_b200.h): every mask value is a class id.  The reference's IoU
        # (models/losses.py:52-60) counts a prediction on a pixel with a foreign label towards the union; the [C,C]
        # confusion tensor has no cell for that, so such a dataset is refused here instead of being scored differently.
        if self.masks.numel() and int(self.masks.max()) >= self.C:
            raise ValueError(f"mask values must be class ids in [0, {self.C}); found {int(self.masks.max())}")
        self.engine = model.engine(self.device, max_microbatch)
        # result buffers: a small ring, so that a level's all-reduce (NCCL's own stream) can still be in flight while
        # the next levels compute into other buffers (submit_level); run_level uses slot 0 synchronously
        self._ring = [(torch.zeros((self.n_batches, self.C, self.C), device=self.device, dtype=torch.int64),
                       torch.zeros((self.n_batches, self.C, 3), device=self.device, dtype=torch.float64))
                      for _ in range(3)]
        self._ring_pending = [None] * len(self._ring)
        self._ring_next = 0
        self.cm, self.dice = self._ring[0]

    def _set_range(self, images_u8, masks_u8, weights):
        self.b0, self.b1 = batch_range(self.n_batches, self.rank, self.world, weights)
        i0, i1 = self.b0 * self.ref_batch, min(self.b1 * self.ref_batch, self.n_total)
        self.i0, self.i1 = i0, i1
        self.images = images_u8[i0:i1].to(self.device, non_blocking=True).contiguous()
        self.masks = masks_u8[i0:i1].to(self.device, non_blocking=True).contiguous()
        for attr in ("_copy_stream", "_back", "_copy_done", "_front_free"):  # upload_async's buffers have the old size
            if hasattr(self, attr):
                delattr(self, attr)

    def rebalance(self, images_u8, masks_u8, rate: float):
        """Re-shard the reference batches in proportion to every rank's measured `rate` (images/s of its own slice,
        e.g. from `local_rate`) and upload the new slice.  Collective: every rank must call it."""
        if self.world == 1:
            return [float(rate)]
        t = torch.zeros((self.world,), device=self.device, dtype=torch.float64)
        t[self.rank] = float(rate)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        rates = [float(v) for v in t.cpu()]
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self._set_range(images_u8, masks_u8, rates)
        return rates

    def local_rate(self, perturbation: Optional[Perturbation], steps: int = 3) -> float:
        """Images/s of THIS rank on its current slice (no collective), by CUDA events: the input of `rebalance`."""
        if self.n_local == 0:
            return 1.0
        self.run_level(perturbation, reduce=False)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            self.run_level(perturbation, reduce=False)
        e1.record()
        e1.synchronize()
        return self.n_local * steps / (e0.elapsed_time(e1) * 1e-3)

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
        # (several micro-batches per pass: the library's ViT and U-Net split it by their own wave quanta, 122 and 148 images)
        step = max(1, 8 * self.engine.max_microbatch // self.ref_batch) * self.ref_batch
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

    def _fill(self, cm, dice, perturbation, want_dice):
        """This rank's rows of one level into (cm, dice); the other ranks' rows are left zero for the sum."""
        if self.world > 1:  # (one rank: the callee zeroes exactly the rows it fills — no extra memset launches)
            cm.zero_()
            if want_dice:
                dice.zero_()
        if self.n_local > 0:
            self._eval(self.images, self.masks, self.local_aux(perturbation), want_dice, cm[self.b0:self.b1],
                       dice[self.b0:self.b1] if want_dice else None)

    def run_level(self, perturbation: Optional[Perturbation], want_dice: bool = False, reduce: bool = True):
        """-> (cm int64 [n_batches,C,C], dice_sums f64 [n_batches,C,3] or None), complete on every rank."""
        if self._ring_pending[0] is not None:
            self._ring_pending[0].wait()
            self._ring_pending[0] = None
        self._fill(self.cm, self.dice, perturbation, want_dice)
        if reduce and self.world > 1:
            dist.all_reduce(self.cm, op=dist.ReduceOp.SUM)
            if want_dice:
                dist.all_reduce(self.dice, op=dist.ReduceOp.SUM)
        return self.cm, (self.dice if want_dice else None)

    def submit_level(self, perturbation: Optional[Perturbation], want_dice: bool = False) -> PendingLevel:
        """Like run_level, but the all-reduce is issued with async_op=True: it runs on NCCL's stream (ordered after
        this level's kernels) and the compute stream goes straight on to the next level, so a rank never idles at a
        level boundary waiting for a slower board.  The result is consumed through `PendingLevel.wait()` — e.g. when
        the level's CSV row is written.  At most len(ring) levels are in flight; an older one is waited for when its
        buffer comes round again."""
        slot = self._ring_next
        self._ring_next = (slot + 1) % len(self._ring)
        if self._ring_pending[slot] is not None:
            self._ring_pending[slot].wait()
        cm, dice = self._ring[slot]
        self._fill(cm, dice, perturbation, want_dice)
        works = []
        if self.world > 1:
            works.append(dist.all_reduce(cm, op=dist.ReduceOp.SUM, async_op=True))
            if want_dice:
                works.append(dist.all_reduce(dice, op=dist.ReduceOp.SUM, async_op=True))
        pending = PendingLevel(cm, dice if want_dice else None, works)
        self._ring_pending[slot] = pending
        return pending

    def run_level_per_batch(self, spec_for_batch, want_dice: bool = False, chunk_batches: int = 16):
        """Like run_level, but the perturbation's random numbers are drawn batch by batch in loader
        order through `spec_for_batch(shape_probe[b,3,H,W]) -> Perturbation` (what the reference's
        nn.Module augmentations do inside the loop, model_wrappers.py:450), `chunk_batches` at a time."""
        if self._ring_pending[0] is not None:
            self._ring_pending[0].wait()
            self._ring_pending[0] = None
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
