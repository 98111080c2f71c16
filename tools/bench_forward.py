"""Forward-only timing of the two models that have no evaluation loop in the reference:
`Autoencoder` (models/autoencoder.py:45-58) and `ClipUnetPrompt` (models/prompt_segmentation.py:31-95, library part:
the CLIP embedding is precomputed, as for ClipUnet's library share).  Inputs are resident fp32 tensors; CUDA events on
the current stream, W warm-up + K timed passes; prints one JSON line per model.  Run on the GPU box:
    python tools/bench_forward.py [--images 1184] [--steps 10] [--warmup 3]
Secondary numbers for DESIGN.md 5b; the headline benchmark is bench.py."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from image_segmentation_b200 import synthetic


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=1184)  # 8 micro-batches of 148
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    n, H, W = args.images, 256, 256
    base, _ = synthetic.make_dataset(148, H, W, seed=99)
    x = synthetic.to_float_nchw(base).repeat(-(-n // 148), 1, 1, 1)[:n].contiguous().to(dev)  # 931 MB at 1184 images: > L2

    from image_segmentation_b200.models.autoencoder import Autoencoder
    torch.manual_seed(42)
    ae = Autoencoder().eval().to(dev)
    eng = ae.engine(dev)
    l0 = eng.launch_count
    ms = timed(lambda: eng.forward(x), args.steps, args.warmup)
    # GEMM work per image: encoder + bottleneck 4 x (1.208 + 2.416) GMAC; decoder 3 x ConvT 0.537 + 3 x 2 x 2.416 (256ch@64^2,
    # 128ch@128^2, 64ch@256^2) = 30.60 GMAC
    gmac = 4 * 3.624 + 3 * 0.537 + 6 * 2.416
    print(json.dumps({"model": "Autoencoder 256x256 forward (reconstruction)", "images_per_step": n, "ms_per_step": ms,
                      "images_per_s": n / ms * 1e3, "gpu_launches_per_step": (eng.launch_count - l0) // (args.steps + args.warmup),
                      "gemm_tflops": 2 * gmac * n / ms, "dtype": "bf16", "inputs": "resident fp32, larger than L2"}), flush=True)
    del ae, eng
    torch.cuda.empty_cache()

    from transformers import CLIPConfig, CLIPModel
    from image_segmentation_b200.models.prompt_segmentation import ClipUnetPrompt
    torch.manual_seed(42)
    pm = ClipUnetPrompt(clip_model=CLIPModel(CLIPConfig()).eval()).eval().to(dev)
    heat = synthetic.prompt_heatmaps(148, H, W, seed=4).repeat(-(-n // 148), 1, 1, 1)[:n].contiguous().to(dev)
    with torch.no_grad():
        feats = torch.cat([pm.clip_feature_extractor(x[i:i + 148]) for i in range(0, n, 148)])
    eng = pm.engine(dev)
    l0 = eng.launch_count
    ms = timed(lambda: eng.forward_prompt(x, heat, feats), args.steps, args.warmup)
    print(json.dumps({"model": "ClipUnetPrompt 256x256 forward, library part (CLIP embedding precomputed)", "images_per_step": n,
                      "ms_per_step": ms, "images_per_s": n / ms * 1e3,
                      "gpu_launches_per_step": (eng.launch_count - l0) // (args.steps + args.warmup), "dtype": "bf16",
                      "inputs": "resident fp32, larger than L2"}), flush=True)


if __name__ == "__main__":
    main()
