#!/usr/bin/env python
"""bench.py — headline benchmark: UNet 256^2 robustness-evaluation images/s (BASELINE.json configs[1]).

One STEP = one (perturbation family, level) pass of the hot path over the 3,680-image synthetic set:
perturb -> U-Net forward -> argmax -> per-batch confusion counts (+ one all-reduce when N > 1).  The
80 (family, level) pairs of scripts/robustness_evaluation.py:59-92 are visited round-robin.

  python bench.py [--gpus N] [--steps K] [--warmup W]            (N > 1: under torchrun, one rank per GPU)
  python bench.py --impl reference ...    times the CPU oracle port of the same path on the host cores

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


def cpu_time_sample(sd, images, masks, levels, n_sample, reps):
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    oracle_step(sd, images[:REF_BATCH], masks[:REF_BATCH], *levels[0], REF_BATCH)  # warm-up
    t0 = time.perf_counter()
    for r in range(reps):
        oracle_step(sd, images[:n_sample], masks[:n_sample], *levels[r % len(levels)], REF_BATCH)
    dt = time.perf_counter() - t0
    return n_sample * reps / dt, cores, dt


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
    checkpoint needs the network) is HuggingFace transformers on PyTorch fp32 kernels, as in the reference, and is
    inside the timed step; the library's share is the UNet minus the (dead) bottleneck block."""
    global CLIP, GFLOP_PER_IMAGE, GEMM_GFLOP_PER_IMAGE
    CLIP = True
    GFLOP_PER_IMAGE = 45.391 - 2 * 3.624          # without bottleneck.conv.0/.3 (1.208 + 2.416 GMAC)
    GEMM_GFLOP_PER_IMAGE = 2 * (21.744 - 3.624 + 0.940)


def net_class():
    from image_segmentation_b200.models.UNet import LargeUNet, UNet
    if CLIP:
        from transformers import CLIPConfig, CLIPModel
        from image_segmentation_b200.models.CLIP_models import ClipUnet
        dt = torch.bfloat16 if os.environ.get("UB_CLIP_DTYPE") == "bf16" else None   # default fp32, as the reference
        return lambda: ClipUnet(clip_model=CLIPModel(CLIPConfig()).eval(), clip_autocast_dtype=dt)
    return LargeUNet if LARGE else UNet


def base_line(args, world):
    if CLIP:
        return {"metric": "ClipUnet 256x256 robustness-eval images/s", "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "data": "synthetic",
                "config": {"workload": "robustness_evaluation sweep levels over 3,680 synthetic 256x256 RGB images, ClipUnet "
                                       "random-init (CLIP ViT-B/32 random-init via transformers, "
                                       + ("bf16 autocast" if os.environ.get("UB_CLIP_DTYPE") == "bf16" else "fp32") + ", inside the step), "
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
    n_sample = 4 if LARGE else 16
    images, masks = synthetic.make_dataset(n_sample, H, W, seed=99)
    levels = level_list()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    for i in range(args.warmup):
        oracle_step(sd, images[:REF_BATCH], masks[:REF_BATCH], *levels[i % 80], REF_BATCH)
    t0 = time.perf_counter()
    for i in range(args.steps):
        oracle_step(sd, images, masks, *levels[i % 80], REF_BATCH)
    dt = time.perf_counter() - t0
    v = n_sample * args.steps / dt
    line = base_line(args, world)
    sample = f"each step = {n_sample} of the {N_IMAGES:,} images of one step (oracle port of the reference path, torch CPU fp32)"
    line.update({"impl": "reference", "value": v, "ms_per_step": dt / args.steps * 1e3, "dtype": "f32",
                 "cpu_baseline": {"value": v, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
                 "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- GPU arm
def run_gpu(args, rank, world, local_rank):
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

    def step(i):
        if LARGE:
            return ev.run_level(None)
        name, lv = levels[i % len(levels)]
        return ev.run_level(make_spec(name, lv, H * W, seed=i))

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

    for i in range(args.warmup):
        step(i)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0 = eng.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = eng.launch_count - l0
    clocks = sampler.stop() if rank == 0 else None
    value = N_IMAGES * args.steps / (ms * 1e-3)

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

    # ---- roofline of the dominant kernel (tcgen05 implicit-GEMM conv / ConvT), per-launch CUDA events
    eng.profile_begin()
    step(1); step(2)
    gemm_ms, gemm_launches, rows = eng.profile_end()
    pk, pk_src = peaks()
    gemm_tflops = GEMM_GFLOP_PER_IMAGE * 1e9 * ev.n_local * 2 / (gemm_ms * 1e-3) / 1e12
    peak = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    roofline = {"bound": "tensor", "kernel": "ub::conv_tc_kernel (all conv3x3 + ConvTranspose launches)",
                "achieved": gemm_tflops, "peak": peak, "unit": "TFLOP/s", "frac": gemm_tflops / peak,
                "peak_source": pk_src + ", sustained figure (kernel timed inside a long step)",
                "frac_of_burst_peak": gemm_tflops / pk["bf16_tflops"], "traffic": None, "traffic_unit": "bytes/launch",
                "gemm_share_of_step": gemm_ms / 2 / (ms / args.steps),
                "flops_per_launch": GEMM_GFLOP_PER_IMAGE * 1e9 * ev.n_local * 2 / max(gemm_launches, 1),
                "avg_launch_ms": gemm_ms / max(gemm_launches, 1)}
    # DRAM bytes per launch from the committed ncu --set full capture (same micro-batch, headline workload only)
    try:
        with open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json")) as f:
            tr = json.load(f)
        if not LARGE and not CLIP:
            per_launch_images = ev.n_local * 2 / max(gemm_launches, 1) * 20   # images per micro-batch in this run
            roofline["traffic"] = tr["dram_bytes_per_image"] * per_launch_images / 20
            roofline["traffic_source"] = (f"profiles/r01_ncu_traffic.json: dram read+write per image from an ncu --set full "
                                          f"capture at micro-batch {tr['microbatch']}, x images per launch here, mean over "
                                          f"the 20 GEMM launches of a micro-batch")
            # every layer's input read once + output written once, bf16 blocked tensors, per image
            roofline["traffic_algorithmic"] = 89.2e6 * per_launch_images / 20   # every GEMM layer's input once + output once (the stem's two readers read the 16-channel packed image)
    except Exception:
        pass
    if args.layers and rank == 0:
        for k, cin, nt, ks, n, t in rows:
            fl = 2.0 * ks * ks * cin * nt
            print(f"# layer {k:28s} cin {cin:4d} n {nt:4d} k{ks} launches {n:4d} ms {t:8.3f}", file=sys.stderr)

    if rank != 0:
        return
    line = base_line(args, world)
    line.update({"value": value, "ms_per_step": ms / args.steps, "dtype": "bf16",
                 "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
                         "steps": k_e2e},
                 "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
                 "whole_net_tflops": value * GFLOP_PER_IMAGE / 1e3, "microbatch": eng.max_microbatch})
    if world == 1 and not args.no_cpu and not CLIP:  # (no CPU leg for the secondary ClipUnet workload)
        lv = [("contrast_increase", 1.1), ("gaussian_blur", 5), ("gaussian_noise", 10), ("occlusion", 25)]
        n_cpu = 8 if LARGE else 64
        v, cores, dt = cpu_time_sample(sd_cpu, images[:n_cpu], masks[:n_cpu], lv, n_cpu, 4)
        line["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": cores, "kind": "port",
                                "sample": f"4 levels (contrast 1.1, blur 5, noise 10, occlusion 25) x {n_cpu} images "
                                          f"through the oracle port, {dt:.1f} s"}
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
