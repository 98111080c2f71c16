"""Drop-in for the reference's scripts/robustness_evaluation.py:27-130: 8 perturbation families x 10
levels, batch 8, soft Dice per batch averaged over batches, one CSV `results/robustness_scores.csv`
with columns perturbation_type,param_value,mean_dice ("%.4f").

Differences forced by this environment, both stated in DESIGN.md: `main` defaults to `UNet` (the reference
script hard-codes ClipUnet, whose CLIP weights need the network; `run_sweep` takes either model and
`main(model_name="ClipUnet")` builds it) and the dataset is the synthetic stand-in unless a dataset object is passed.  Random families use the device RNG unless
`parity_rng=True`, in which case the reference's host draws are replayed (slow for salt-and-pepper,
as in the reference).  Multi-GPU: launch under torchrun; batches are sharded over ranks and rank 0
writes the CSV.
"""
from __future__ import annotations

import csv
import os
import random

import numpy as np
import torch
import torch.distributed as dist

from ..engine import Perturbation
from ..evaluator import RobustnessEvaluator

PERTURBATIONS = {  # robustness_evaluation.py:59-92, same order
    "gaussian_noise": [0, 2, 4, 6, 8, 10, 12, 14, 16, 18],
    "gaussian_blur": [0, 1, 2, 3, 4, 5, 6, 7, 8, 9],
    "contrast_increase": [1.0, 1.01, 1.02, 1.03, 1.04, 1.05, 1.1, 1.15, 1.2, 1.25],
    "contrast_decrease": [1.0, 0.95, 0.9, 0.85, 0.8, 0.6, 0.4, 0.3, 0.2, 0.1],
    "brightness_increase": [0, 5, 10, 15, 20, 25, 30, 35, 40, 45],
    "brightness_decrease": [0, 5, 10, 15, 20, 25, 30, 35, 40, 45],
    "occlusion": [0, 5, 10, 15, 20, 25, 30, 35, 40, 45],
    "salt_pepper_noise": [0.00, 0.02, 0.04, 0.06, 0.08, 0.10, 0.12, 0.14, 0.16, 0.18],
}


def make_spec(name: str, level, hw: int, seed: int = 0) -> Perturbation:
    if name == "gaussian_noise":
        return Perturbation.u8_noise(level, seed=seed)
    if name == "gaussian_blur":
        return Perturbation.u8_blur(level)
    if name in ("contrast_increase", "contrast_decrease"):
        return Perturbation.u8_contrast(level)
    if name == "brightness_increase":
        return Perturbation.u8_brightness(level, True)
    if name == "brightness_decrease":
        return Perturbation.u8_brightness(level, False)
    if name == "occlusion":
        return Perturbation.u8_occlusion(level, seed=seed)
    if name == "salt_pepper_noise":
        return Perturbation.u8_salt_pepper(level, seed=seed, hw=hw)
    raise KeyError(name)


def parity_spec(name: str, level, n: int, h: int, w: int) -> Perturbation:
    """Replays the reference's per-sample host draws, in dataset order (one pass of its DataLoader)."""
    from ..customDatasets import perturbations as P
    spec = make_spec(name, level, h * w)
    if name == "gaussian_noise":
        spec.aux = torch.stack([torch.normal(mean=0.0, std=float(level), size=(3, h, w), dtype=torch.float)
                                for _ in range(n)])
    elif name == "occlusion":
        ds = P.OcclusionIncrease(None, level)
        spec.aux = torch.tensor([ds.draw_corner(h, w) for _ in range(n)], dtype=torch.int32).reshape(n, 2)
    elif name == "salt_pepper_noise":
        ds = P.SaltPepperNoise(None, level)
        spec.aux = torch.from_numpy(np.stack([ds.draw_code(h, w) for _ in range(n)]))
    return spec


def run_sweep(model, images_u8, masks_u8, results_file="results/robustness_scores.csv", batch_size=8,
              parity_rng=False, perturbations=None, device=None, quiet=False):
    """-> list of (perturbation_type, param_value, mean_dice); rank 0 writes the CSV."""
    random.seed(42); np.random.seed(42); torch.manual_seed(42)  # robustness_evaluation.py:32-34
    ev = RobustnessEvaluator(model, images_u8, masks_u8, batch_size, device)
    n, h, w = images_u8.shape[:3]
    rank0 = ev.rank == 0
    rows = []
    f = writer = None
    if rank0:
        os.makedirs(os.path.dirname(results_file) or ".", exist_ok=True)
        f = open(results_file, "w", newline="")
        writer = csv.writer(f)
        writer.writerow(["perturbation_type", "param_value", "mean_dice"])

    def consume(item):
        # the level's all-reduce (issued async on NCCL's stream) is waited for only here, when its CSV row is written
        p_name, level, pending = item
        _, dice = pending.wait()
        mean_dice = RobustnessEvaluator.script_score(dice)
        rows.append((p_name, level, mean_dice))
        if rank0:
            writer.writerow([p_name, level, f"{mean_dice:.4f}"])
            if not quiet:
                print(f"{p_name} param={level}, Dice={mean_dice:.4f}")

    try:
        in_flight = None   # one level computes while the previous one's row is formatted: the GPU never waits for the host
        for li, (p_name, levels) in enumerate((perturbations or PERTURBATIONS).items()):
            for lj, level in enumerate(levels):
                spec = parity_spec(p_name, level, n, h, w) if parity_rng else make_spec(p_name, level, h * w, seed=1000 * li + lj)
                pending = ev.submit_level(spec, want_dice=True)
                if in_flight is not None:
                    consume(in_flight)
                in_flight = (p_name, level, pending)
        if in_flight is not None:
            consume(in_flight)
    finally:
        if f:
            f.close()
    return rows


def main(n_images=3680, checkpoint="saved-models/UNet/model.pth", model_name="UNet", clip_model=None):
    """model_name "UNet" (BASELINE.json's metric) or "ClipUnet" (what the reference script hard-codes, :36-42; needs the
    openai/clip-vit-base-patch32 checkpoint in the HuggingFace cache, or a CLIPModel passed as `clip_model`)."""
    from ..models.UNet import UNet
    from ..synthetic import make_dataset
    if "RANK" in os.environ and not dist.is_initialized():
        dist.init_process_group("nccl")
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    device = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(42)
    if model_name == "ClipUnet":
        from ..models.CLIP_models import ClipUnet
        model = ClipUnet(clip_model=clip_model)
    else:
        model = UNet()
    if os.path.exists(checkpoint):
        ckpt = torch.load(checkpoint, map_location="cpu")
        model.load_state_dict({k.replace("_orig_mod.", ""): v for k, v in ckpt.items()}, strict=False)
    model = model.to(device).eval()
    images, masks = make_dataset(n_images)
    run_sweep(model, images, masks, device=device)
    if (not dist.is_initialized()) or dist.get_rank() == 0:
        print("Evaluation complete. Results saved to CSV.")


if __name__ == "__main__":
    main()
