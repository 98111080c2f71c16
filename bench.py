#!/usr/bin/env python
"""bench.py — headline benchmark: UNet 256^2 robustness-evaluation images/s (BASELINE.json configs[1]).

One STEP = one (perturbation family, level) pass of the hot path over the 3,680-image synthetic set:
perturb -> U-Net forward -> argmax -> per-batch confusion counts (+ one all-reduce when N > 1).  The
80 (family, level) pairs of scripts/robustness_evaluation.py:59-92 are visited round-robin.

  python bench.py [--gpus N] [--steps K] [--warmup W]            (N > 1: under torchrun, one rank per GPU)
  python bench.py --impl reference ...    times the REFERENCE'S OWN modules (baseline/_ref: its Dataset-wrapper
                                          perturbations, UNet.forward, IoU / PixelAccuracy; byte-for-byte copies made by
                                          baseline/install_ref.py) on the host cores; the oracle port only when that
                                          directory is absent

Prints ONE JSON line on rank 0 (contract in the task description): value = device-resident throughput,
e2e = through the public API with pinned HOST buffers (H2D of the step's images+masks and D2H of the
confusion tensor inside the timed region), roofline = tensor-core GEMM kernels vs MEASURED_PEAKS.json,
cpu_baseline = the oracle on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

N_IMAGES = 3680          # Oxford-IIIT-Pet test split size the reference evaluates
REF_BATCH = 8            # scripts/robustness_evaluation.py:112
H = W = 256
GFLOP_PER_IMAGE = 45.391  # all Conv2d + ConvTranspose2d MACs x2 of UNet@256^2
GEMM_GFLOP_PER_IMAGE = 2 * (21.744 + 0.940)  # conv3x3 + ConvT only (the tensor-core kernel), SURVEY 8a


def level_list():
    from image_segmentation_b200.scripts.robustness_evaluation import PERTURBATIONS
    pairs = [(name, lv) for name, lvls in PERTURBATIONS.items() for lv in lvls]
    return [pairs[(i * 11) % len(pairs)] for i in range(len(pairs))]  # interleave the families


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return p, "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True).start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU arm (oracle port)
def oracle_step(sd, images_u8_nhwc, masks_u8, name, level, ref_batch):
    """The reference path restated on the CPU (oracle/): perturb -> fp32 forward -> metrics per batch."""
    import random
    import numpy as np
    from oracle import metrics_oracle as mo, perturb_oracle as po, unet_oracle
    u8 = images_u8_nhwc.permute(0, 3, 1, 2).numpy()
    n, _, h, w = u8.shape
    if name == "gaussian_noise":
        out = np.stack([po.u8_gaussian_noise(u8[i], torch.normal(0.0, float(level), size=(3, h, w)).numpy()) for i in range(n)])
    elif name == "gaussian_blur":
        out = po.u8_blur(u8, level)
    elif name.startswith("contrast"):
        out = po.u8_contrast(u8, level)
    elif name == "brightness_increase":
        out = po.u8_brightness(u8, level, True)
    elif name == "brightness_decrease":
        out = po.u8_brightness(u8, level, False)
    elif name == "occlusion":
        out = np.stack([po.u8_occlusion(u8[i], level, *po.replay_occlusion_corner(level, h, w, random)) for i in range(n)])
    else:
        out = np.stack([po.u8_salt_pepper(u8[i], po.replay_salt_pepper_code(level, h, w, random)) for i in range(n)])
    x = torch.from_numpy(po.dequantise(out))
    scores = []
    for b0 in range(0, n, ref_batch):
        logits = unet_oracle.forward(sd, x[b0:b0 + ref_batch])
        t = masks_u8[b0:b0 + ref_batch]
        cm = mo.confusion_matrix(unet_oracle.hard_prediction(logits).numpy(), t.numpy())
        scores.append((mo.iou_from_cm(cm), mo.pixel_accuracy_from_cm(cm), mo.dice_smp(logits, t.long())))
    return scores


REF_CLASS = {"gaussian_noise": "GaussianPixelNoise", "gaussian_blur": "GaussianBlur", "contrast_increase": "ContrastIncrease",
             "contrast_decrease": "ContrastDecrease", "brightness_increase": "BrightnessIncrease",
             "brightness_decrease": "BrightnessDecrease", "occlusion": "OcclusionIncrease",
             "salt_pepper_noise": "SaltPepperNoise"}   # scripts/robustness_evaluation.py:59-92


def load_reference():
    """The reference's own modules from baseline/_ref (None when the directory did not travel)."""
    root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(root, "models", "UNet.py")):
        return None
    from oracle import refload
    return refload.load(root)


class _RefBase(torch.utils.data.Dataset):
    """Item contract of customDatasets/datasets.py:137-160: (fp32 [3,H,W] = u8/255, int64 [H,W])."""

    def __init__(self, images_u8_nhwc, masks_u8):
        self.x = images_u8_nhwc.permute(0, 3, 1, 2).float() / 255.0
        self.t = masks_u8.long()

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, i):
        return self.x[i], self.t[i]


def reference_step(ns, model, images_u8_nhwc, masks_u8, name, level, ref_batch):
    """One (family, level) pass exactly as the reference runs it (scripts/robustness_evaluation.py:105-124 with the
    metric block of models/model_wrappers.py:456-458): its Dataset wrapper around the base set, a DataLoader of
    `ref_batch`, UNet.forward under no_grad, IoU / PixelAccuracy / 2I/(1+I) per batch — all the reference's own code —
    plus the restated smp Dice (the package is absent; oracle/metrics_oracle.py)."""
    from oracle import metrics_oracle as mo
    ds = getattr(ns.perturbations, REF_CLASS[name])(_RefBase(images_u8_nhwc, masks_u8), level)
    loader = torch.utils.data.DataLoader(ds, batch_size=ref_batch, shuffle=False)
    iou_fn, pa_fn = ns.losses.IoU(), ns.losses.PixelAccuracy()
    tot = [0.0, 0.0, 0.0, 0.0]
    for x, t in loader:
        with torch.no_grad():
            z = model(x)
        iou = iou_fn(z, t)
        tot[0] += iou.item(); tot[1] += pa_fn(z, t).item(); tot[2] += (2 * iou / (1 + iou)).item()
        tot[3] += mo.dice_smp(z, t)
    return [v / len(loader) for v in tot]


def make_cpu_arm(sd):
    """-> (step(images, masks, name, level), kind, description)"""
    ns = load_reference()
    if ns is not None and not LARGE:
        model = ns.UNet().eval()
        model.load_state_dict(sd)
        return (lambda im, mk, name, lv: reference_step(ns, model, im, mk, name, lv, REF_BATCH), "reference",
                "the reference's own modules (baseline/_ref: customDatasets/perturbations.py wrappers -> DataLoader -> "
                "models/UNet.py forward fp32 -> models/losses.py IoU + PixelAccuracy, restated smp Dice)")
    if ns is not None and LARGE:
        model = ns.LargeUNet().eval()
        model.load_state_dict(sd)

        def step(im, mk, name, lv):
            with torch.no_grad():
                z = model(im.permute(0, 3, 1, 2).float() / 255.0)
            return [ns.losses.IoU()(z, mk.long()).item(), ns.losses.PixelAccuracy()(z, mk.long()).item()]
        return step, "reference", "the reference's own LargeUNet.forward + IoU + PixelAccuracy (baseline/_ref), fp32"
    return (lambda im, mk, name, lv: oracle_step(sd, im, mk, name, lv, REF_BATCH), "port",
            "oracle port of the reference path (baseline/_ref absent), torch CPU fp32")


def cpu_time_sample(sd, images, masks, levels, n_sample, reps):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step, kind, what = make_cpu_arm(sd)
    step(images[:REF_BATCH], masks[:REF_BATCH], *levels[0])  # warm-up
    t0 = time.perf_counter()
    for r in range(reps):
        step(images[:n_sample], masks[:n_sample], *levels[r % len(levels)])
    dt = time.perf_counter() - t0
    return n_sample * reps / dt, cores, dt, kind, what


CLIP = False             # --workload clipunet256: the model the reference's sweep script hard-codes (secondary)
LARGE = False            # --workload large512: BASELINE.json configs[2] (LargeUNet, 512x512, 64 images per GPU)


def use_large_workload(world):
    """Secondary workload: LargeUNet @512^2, 64 images per GPU per step, no perturbation
    (forward + metrics).  Not the headline; the default run is unchanged."""
    global LARGE, N_IMAGES, REF_BATCH, H, W, GFLOP_PER_IMAGE, GEMM_GFLOP_PER_IMAGE
    LARGE = True
    N_IMAGES, REF_BATCH, H, W = 64 * world, 64, 512, 512
    GFLOP_PER_IMAGE = 226.123
    GEMM_GFLOP_PER_IMAGE = 226.123 - 2 * (H * W * 3 * 32 * 2) / 1e9  # minus the two 1x1 convs


def use_clip_workload():
    """Secondary workload: ClipUnet on the headline sweep.  The CLIP ViT-B/32 embedding (random-init here: the
    checkpoint needs the network) is inside the timed step: the library's own ViT by default, HuggingFace transformers
    on PyTorch kernels with UB_CLIP_BACKEND=transformers[-bf16]; the U-Net part is the UNet minus the (dead) bottleneck."""
    global CLIP, GFLOP_PER_IMAGE, GEMM_GFLOP_PER_IMAGE
    CLIP = True
    GFLOP_PER_IMAGE = 45.391 - 2 * 3.624          # without bottleneck.conv.0/.3 (1.208 + 2.416 GMAC)
    GEMM_GFLOP_PER_IMAGE = 2 * (21.744 - 3.624 + 0.940)


def clip_backend():
    be = os.environ.get("UB_CLIP_BACKEND", "b200")
    if be not in ("b200", "transformers", "transformers-bf16"):
        raise SystemExit(f"UB_CLIP_BACKEND={be!r}: b200 | transformers | transformers-bf16")
    return be


def net_class():
    from image_segmentation_b200.models.UNet import LargeUNet, UNet
    if CLIP:
        from transformers import CLIPConfig, CLIPModel
        from image_segmentation_b200.models.CLIP_models import ClipUnet
        # UB_CLIP_BACKEND: "b200" (default: the library's ViT), "transformers" (HuggingFace on PyTorch kernels, fp32 as the
        # reference) or "transformers-bf16" (the same under bf16 autocast)
        be = clip_backend()
        return lambda: ClipUnet(clip_model=CLIPModel(CLIPConfig()).eval(),
                                clip_autocast_dtype=torch.bfloat16 if be == "transformers-bf16" else None,
                                clip_backend="b200" if be == "b200" else "transformers")
    return LargeUNet if LARGE else UNet


def base_line(args, world):
    if CLIP:
        return {"metric": "ClipUnet 256x256 robustness-eval images/s", "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "data": "synthetic",
                "config": {"workload": "robustness_evaluation sweep levels over 3,680 synthetic 256x256 RGB images, ClipUnet "
                                       "random-init (CLIP ViT-B/32 random-init, "
                                       + {"b200": "the library's ViT (bf16 GEMMs)", "transformers": "transformers fp32",
                                          "transformers-bf16": "transformers under bf16 autocast"}[clip_backend()] + ", inside the step), "
                                       "reference batch 8", "images_per_step": N_IMAGES, "ref_batch": REF_BATCH,
                           "l2": "inputs_exceed_l2 (965 MB of images+masks per step)"}}
    if LARGE:
        return {"metric": "LargeUNet 512x512 forward+metrics images/s", "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "data": "synthetic",
                "config": {"workload": "LargeUNet random-init (seed 42), 64 synthetic 512x512 RGB images per GPU per step, "
                                       "3 classes, forward + argmax + confusion counts, no perturbation",
                           "images_per_step": N_IMAGES, "ref_batch": REF_BATCH,
                           "l2": "activations_exceed_l2 (every layer's tensors are > 126 MB at 64 images)"}}
    return {"metric": "UNet 256x256 robustness-eval images/s", "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "data": "synthetic",
            "config": {"workload": "robustness_evaluation sweep levels (8 perturbations x 10 levels, visited round-robin) over "
                                   "3,680 synthetic 256x256 RGB images, 3 classes, UNet random-init (seed 42), reference batch 8",
                       "images_per_step": N_IMAGES, "ref_batch": REF_BATCH, "l2": "inputs_exceed_l2 (965 MB of images+masks per step)"}}


def run_reference(args, rank, world):
    if rank != 0:
        return
    from image_segmentation_b200 import synthetic
    torch.manual_seed(42)
    sd = net_class()().eval().state_dict()
    n_sample = 4 if LARGE else 64
    images, masks = synthetic.make_dataset(n_sample, H, W, seed=99)
    levels = level_list()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    step, kind, what = make_cpu_arm(sd)
    for i in range(args.warmup):
        step(images[:REF_BATCH], masks[:REF_BATCH], *levels[i % 80])
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(images, masks, *levels[(args.warmup + i) % 80])
    dt = time.perf_counter() - t0
    v = n_sample * args.steps / dt
    line = base_line(args, world)
    sample = f"each step = {n_sample} of the {N_IMAGES:,} images of one (family, level) pass; {what}; {cores} threads"
    line.update({"impl": "reference", "value": v, "ms_per_step": dt / args.steps * 1e3, "dtype": "f32",
                 "cpu_baseline": {"value": v, "unit": "images/s", "cores": cores, "kind": kind, "sample": sample},
                 "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU arm
def timed_forward(fn, warmup=3, steps=10):
    """ms per call of fn() by CUDA events on the current stream (secondary measurements)."""
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / steps


def extra_measurements(dev):
    """Short secondary measurements for the rows after the headline, run AFTER the headline's timed
    region on the same GPU: each is a few hundred ms of device time.  Synthetic inputs, random-init weights (values do
    not affect throughput); inputs resident on the device; images/s = images per call / mean call time."""
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.evaluator import RobustnessEvaluator
    out = {}
    # (f1) BASELINE.json configs[2]: LargeUNet, 512x512, 64 images per GPU, forward + argmax + confusion counts
    try:
        from image_segmentation_b200.models.UNet import LargeUNet
        torch.manual_seed(42)
        m = LargeUNet().eval().to(dev)
        bi, bm = synthetic.make_dataset(8, 512, 512, seed=3)
        ev = RobustnessEvaluator(m, bi.repeat(8, 1, 1, 1), bm.repeat(8, 1, 1), 64, dev, rank=0, world_size=1, max_microbatch=64)
        ms = timed_forward(lambda: ev.run_level(None))
        out["large512"] = {"images_per_s": 64 / (ms * 1e-3), "ms_per_step": ms, "images_per_step": 64,
                           "tflops": 64 * 226.123e9 / (ms * 1e-3) / 1e12,
                           "what": "LargeUNet 512x512, 64 images: forward + argmax + counts (configs[2])"}
        del ev, m
    except Exception as e:  # a secondary line must never take the headline down
        out["large512"] = {"error": repr(e)[:200]}
    # (f2) configs[3]: SegmentationModel(Encoder, SegmentationDecoderSkip) at 256x256, 128 images per GPU
    try:
        from image_segmentation_b200.models.autoencoder import Encoder
        from image_segmentation_b200.models.pre_trained import SegmentationDecoderSkip, SegmentationModel
        torch.manual_seed(42)
        m = SegmentationModel(Encoder(), SegmentationDecoderSkip(3)).eval().to(dev)
        bi, bm = synthetic.make_dataset(16, 256, 256, seed=4)
        ev = RobustnessEvaluator(m, bi.repeat(8, 1, 1, 1), bm.repeat(8, 1, 1), 8, dev, rank=0, world_size=1, max_microbatch=128)
        ms = timed_forward(lambda: ev.run_level(None))
        out["segmodel_b128"] = {"images_per_s": 128 / (ms * 1e-3), "ms_per_step": ms, "images_per_step": 128,
                                "what": "SegmentationModel (autoencoder Encoder + skip decoder) 256x256, 128 images: "
                                        "forward + argmax + counts (configs[3])"}
        del ev, m
    except Exception as e:
        out["segmodel_b128"] = {"error": repr(e)[:200]}
    # (f4) configs[4]: ClipUnetPrompt decoder over frozen CLIP features with the prompt heat-map channel at 224x224; the
    # CLIP embedding is an INPUT here (precomputed, "frozen features"): the library part only
    try:
        from image_segmentation_b200.models.prompt_segmentation import ClipUnetPrompt
        torch.manual_seed(42)
        m = ClipUnetPrompt(clip_model=torch.nn.Identity()).eval().to(dev)   # (the CLIP encoder is not called)
        eng = m.engine(dev, 128)
        x = synthetic.to_float_nchw(synthetic.make_dataset(16, 224, 224, seed=5)[0]).repeat(8, 1, 1, 1).to(dev)
        heat = synthetic.prompt_heatmaps(128, 224, 224, seed=6).to(dev)
        feats = torch.randn((128, 512), device=dev)
        ms = timed_forward(lambda: eng.forward_prompt(x, heat, feats))
        out["clip_prompt_224"] = {"images_per_s": 128 / (ms * 1e-3), "ms_per_step": ms, "images_per_step": 128,
                                  "what": "ClipUnetPrompt 224x224, 128 images, CLIP embeddings given: stem + enc1-3 + "
                                          "PromptEncoder + fusion + dec1-4 + head -> logits (configs[4])"}
    except Exception as e:
        out["clip_prompt_224"] = {"error": repr(e)[:200]}
    # (f4) the CLIP image embedding itself (models/processing_blocks.py:156-233: Resize + Normalize + ViT-B/32 + projection)
    # in the library, 244 images (two internal passes of 122) at 256x256, random-init tower
    try:
        from transformers import CLIPConfig, CLIPModel
        from image_segmentation_b200.models.processing_blocks import ClipFeatureExtractor
        torch.manual_seed(42)
        ext = ClipFeatureExtractor(clip_model=CLIPModel(CLIPConfig()).eval())
        x = synthetic.to_float_nchw(synthetic.make_dataset(61, 256, 256, seed=7)[0]).repeat(4, 1, 1, 1).to(dev)
        ms = timed_forward(lambda: ext(x))
        out["clip_vit_b32_256"] = {"images_per_s": 244 / (ms * 1e-3), "ms_per_step": ms, "images_per_step": 244,
                                   "tflops": 244 * 8.74e9 / (ms * 1e-3) / 1e12,
                                   "what": "CLIP ViT-B/32 image embedding in the library (resize 256->224 + normalise + "
                                           "12 layers + projection), 244 images"}
        del ext
    except Exception as e:
        out["clip_vit_b32_256"] = {"error": repr(e)[:200]}
    return out


def run_gpu(args, rank, world, local_rank):
    import hashlib
    import torch.distributed as dist
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.evaluator import RobustnessEvaluator
    from image_segmentation_b200.scripts.robustness_evaluation import make_spec
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    torch.manual_seed(42)
    model = net_class()().eval()
    sd_cpu = {k: v.clone() for k, v in model.state_dict().items()}
    model = model.to(dev)
    n_base = min(N_IMAGES, 64 if LARGE else 230)
    base_i, base_m = synthetic.make_dataset(n_base, H, W, seed=99)
    reps = -(-N_IMAGES // n_base)
    images = base_i.repeat(reps, 1, 1, 1)[:N_IMAGES].contiguous().pin_memory()   # values do not affect throughput
    masks = base_m.repeat(reps, 1, 1)[:N_IMAGES].contiguous().pin_memory()
    ev = RobustnessEvaluator(model, images, masks, REF_BATCH, dev, rank=rank, world_size=world,
                             max_microbatch=args.microbatch)
    eng = ev.engine
    levels = level_list()

    def spec_of(i):
        if LARGE:
            return None
        name, lv = levels[i % len(levels)]
        return make_spec(name, lv, H * W, seed=i)

    def step(i):                      # synchronous form: result complete on the compute stream when it returns
        return ev.run_level(spec_of(i))

    def submit(i):                    # the all-reduce of level i stays in flight on NCCL's stream (evaluator.submit_level)
        return ev.submit_level(spec_of(i))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def gather(x):
        if world > 1:
            t = torch.zeros((world,), device=dev, dtype=torch.float64)
            t[rank] = x
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            return [float(v) for v in t.cpu()]
        return [x]

    for i in range(args.warmup):
        submit(i).wait()
    # N > 1: the boards of a box settle at different power-capped clocks and a sharded level ends with its slowest
    # rank, so the reference batches are re-split once, in proportion to each rank's warm rate (untimed, part of warm-up)
    rates = None
    if world > 1 and args.rebalance:
        rates = ev.rebalance(images, masks, ev.local_rate(spec_of(0), steps=2))
        for i in range(2):
            submit(i).wait()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = eng.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    pend = [submit(args.warmup + i) for i in range(args.steps)]
    for p in pend:                     # every level's result is complete on the compute stream before the clock stops
        p.wait()
    e1.record()
    barrier()
    my_ms = e0.elapsed_time(e1)
    ms = max_over_ranks(my_ms)
    per_rank_ms = gather(my_ms)
    launches = eng.launch_count - l0
    clocks = sampler.stop() if rank == 0 else None
    value = N_IMAGES * args.steps / (ms * 1e-3)

    # ---- checksum of one fixed deterministic level after the all-reduce: the same at every N iff N ranks reproduce
    # the single-GPU counts (sharding, micro-batch split and re-balancing must not change a single count)
    cm_sha = None
    if not LARGE and not CLIP:
        cm_fixed, _ = ev.run_level(make_spec("contrast_increase", 1.1, H * W))
        cm_sha = hashlib.sha256(cm_fixed.cpu().numpy().tobytes()).hexdigest()
        assert int(cm_fixed.sum()) == N_IMAGES * H * W or os.environ.get("UB_DBG")   # (UB_DBG ablations produce wrong counts on purpose)

    # ---- end to end through the public API: pinned host buffers -> H2D -> level -> D2H of the counts
    k_e2e = max(1, args.steps)   # the first upload has nothing to overlap with (pipeline fill): amortised over K steps
    h2d = ev.n_local * (H * W * 3 + H * W)
    d2h = ev.cm.numel() * 8
    # warm-up of this path too (untimed): the first upload_async allocates the back buffers and creates the copy stream
    ev.upload_async(images, masks); ev.swap(); step(0); ev.cm.cpu()
    barrier()
    t0 = time.perf_counter()
    ev.upload_async(images, masks)              # step 0's inputs: nothing to overlap with yet
    for i in range(k_e2e):
        ev.swap()                               # compute stream waits for this step's copy
        if i + 1 < k_e2e:
            ev.upload_async(images, masks)      # next step's H2D overlaps this step's kernels (copy stream)
        cm, _ = step(args.warmup + i)
        cm_host = cm.cpu()                      # D2H of the step's result (synchronises)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = N_IMAGES * k_e2e / e2e_s
    h2d_all = sum(int(v) for v in gather(float(h2d)))

    # ---- the collective alone (N > 1): mean duration of the int64 [n_batches,C,C] all-reduce, back to back, CUDA events
    allreduce_us = None
    if world > 1:
        for _ in range(3):
            dist.all_reduce(ev.cm)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a0.record()
        for _ in range(20):
            dist.all_reduce(ev.cm)
        a1.record()
        a1.synchronize()
        allreduce_us = max_over_ranks(a0.elapsed_time(a1) / 20 * 1e3)

    # ---- roofline of the dominant kernel (tcgen05 implicit-GEMM conv / ConvT), per-launch CUDA events
    # (the share of the step is taken against the duration of the SAME two profiled steps: with an event pair around every
    # launch they run without launch overlap, so dividing by the un-profiled step time overstates it — 1.04 at N = 8)
    eng.profile_begin()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    step(1); step(2)
    p1.record()
    gemm_ms, gemm_launches, rows = eng.profile_end()
    prof_ms = p0.elapsed_time(p1)
    pk, pk_src = peaks()
    gemm_tflops = GEMM_GFLOP_PER_IMAGE * 1e9 * ev.n_local * 2 / (gemm_ms * 1e-3) / 1e12
    peak = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    roofline = {"bound": "tensor", "kernel": "ub::conv_tc_kernel (all conv3x3 + ConvTranspose launches)",
                "achieved": gemm_tflops, "peak": peak, "unit": "TFLOP/s", "frac": gemm_tflops / peak,
                "peak_source": pk_src + ", sustained figure (kernel timed inside a long step)",
                "frac_of_burst_peak": gemm_tflops / pk["bf16_tflops"], "traffic": None, "traffic_unit": "bytes/launch",
                "gemm_share_of_step": gemm_ms / prof_ms, "profiled_step_ms": prof_ms / 2,
                "flops_per_launch": GEMM_GFLOP_PER_IMAGE * 1e9 * ev.n_local * 2 / max(gemm_launches, 1),
                "avg_launch_ms": gemm_ms / max(gemm_launches, 1)}
    # DRAM bytes per launch: ncu --set full capture of the same build (profiles/), headline workload only — ncu cannot
    # run inside a timed bench, so this is the capture's per-image figure x the images per launch of THIS run
    try:
        tr_path = next(p for p in (os.path.join(ROOT, "profiles", f) for f in ("r02_ncu_traffic_final.json", "r02_ncu_traffic.json", "r01_ncu_traffic.json"))
                       if os.path.exists(p))
        with open(tr_path) as f:
            tr = json.load(f)
        if not LARGE and not CLIP:
            n_gemm = tr.get("gemm_launches_per_microbatch", 20)
            per_launch_images = ev.n_local * 2 / max(gemm_launches, 1) * n_gemm   # images per micro-batch in this run
            roofline["traffic"] = tr["dram_bytes_per_image"] * per_launch_images / n_gemm
            roofline["traffic_source"] = (f"{os.path.relpath(tr_path, ROOT)}: dram read+write per image from an ncu --set full "
                                          f"capture at micro-batch {tr['microbatch']}, x images per launch here, mean over "
                                          f"the {n_gemm} GEMM launches of a micro-batch")
            # every layer's input read once + output written once, bf16 blocked tensors, per image
            roofline["traffic_algorithmic"] = tr.get("algorithmic_bytes_per_image", 89.2e6) * per_launch_images / n_gemm
    except Exception:
        pass
    if args.layers and rank == 0:
        for k, cin, nt, ks, n, t in rows:
            print(f"# layer {k:28s} cin {cin:4d} n {nt:4d} k{ks} launches {n:4d} ms {t:8.3f}", file=sys.stderr)

    extra = None
    if world == 1 and not args.no_extra and not LARGE and not CLIP:
        del ev
        extra = extra_measurements(dev)

    if rank != 0:
        return
    line = base_line(args, world)
    line.update({"value": value, "ms_per_step": ms / args.steps, "dtype": "bf16",
                 "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h * world,
                         "steps": k_e2e},
                 "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
                 "whole_net_tflops": value * GFLOP_PER_IMAGE / 1e3, "microbatch": eng.max_microbatch,
                 "cm_sha256": cm_sha})
    if world > 1:
        line["multi_gpu"] = {"per_rank_ms_per_step": [v / args.steps for v in per_rank_ms], "allreduce_us": allreduce_us,
                             "warm_rates_images_per_s": rates, "rebalanced": rates is not None,
                             "collective": "one async int64 all-reduce of [n_batches,3,3] per level on NCCL's stream"}
    if extra is not None:
        line["extra"] = extra
    if world == 1 and not args.no_cpu and not CLIP:  # (no CPU leg for the secondary ClipUnet workload)
        lv = [("contrast_increase", 1.1), ("gaussian_blur", 5), ("gaussian_noise", 10), ("occlusion", 25)]
        n_cpu = 8 if LARGE else 64
        v, cores, dt, kind, what = cpu_time_sample(sd_cpu, images[:n_cpu], masks[:n_cpu], lv, n_cpu, 4)
        line["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": cores, "kind": kind,
                                "sample": f"4 levels (contrast 1.1, blur 5, noise 10, occlusion 25) x {n_cpu} images "
                                          f"through {what}, {dt:.1f} s"}
    else:
        line["cpu_baseline"] = None
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--microbatch", type=int, default=None)
    ap.add_argument("--layers", action="store_true", help="print the per-layer GEMM times to stderr")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the short secondary measurements (extra)")
    ap.add_argument("--rebalance", action="store_true",
                    help="N > 1: re-split the reference batches in proportion to each rank's warm rate (default: equal split — on "
                         "power-capped boards of one box the 2-step warm rates are noisier than the boards differ; measured "
                         "-0.5 to -0.7 %% at N = 8)")
    ap.add_argument("--no-rebalance", action="store_true", help="(accepted for older command lines; the equal split is the default)")
    ap.add_argument("--workload", default="unet256", choices=["unet256", "large512", "clipunet256"],
                    help="unet256 = the headline (BASELINE.json configs[1]); large512 = configs[2]; clipunet256: secondary")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.workload == "large512":
        use_large_workload(world)
    if args.workload == "clipunet256":
        use_clip_workload()
    if args.impl == "reference":
        if CLIP:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "no CPU arm for the secondary clipunet256 workload"}))
            return
        run_reference(args, rank, world)
        return
    import torch.distributed as dist
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # rank 0 prints ONE JSON line on stdout: send NCCL's version banner / debug output to stderr instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION", "WARN"):
            os.environ.pop("NCCL_DEBUG", None)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_gpu(args, rank, world, local_rank)
    finally:
        if world > 1:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
