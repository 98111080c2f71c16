"""Times ub_clip_image_features (the library's CLIP ViT-B/32 image tower) on one GPU: images/s for a batch of 128 at 256x256,
next to HuggingFace transformers on the same GPU in fp32 and under bf16 autocast.
Usage: python tools/time_clip_vit.py [n] [ours]     ("ours": one warm-up and one timed pass of the library only — the ncu launch list)"""
import sys
import time

import torch

sys.path.insert(0, ".")
from oracle import clip_oracle as co  # noqa: E402  (seeded weights only; nothing here is timed through the oracle)


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    from image_segmentation_b200.models.processing_blocks import ClipFeatureExtractor
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    dev = torch.device("cuda:0")
    m = co.seeded_clip(seed=3)
    x = torch.rand((n, 3, 256, 256), device=dev)
    ours = ClipFeatureExtractor(clip_model=m)
    hf32 = ClipFeatureExtractor(clip_model=m, backend="transformers")
    hf16 = ClipFeatureExtractor(clip_model=m, backend="transformers", autocast_dtype=torch.bfloat16)
    only_ours = len(sys.argv) > 2 and sys.argv[2] == "ours"
    arms = (("library ViT (bf16 GEMMs)", ours),) if only_ours else (
        ("library ViT (bf16 GEMMs)", ours), ("transformers fp32", hf32), ("transformers bf16 autocast", hf16))
    for name, f in arms:
        ms = timed(lambda: f(x), reps=1 if only_ours else 10, warm=1 if only_ours else 3)
        print(f"{name:32s} {ms:8.3f} ms per {n} images = {n / ms * 1e3:9.0f} images/s")
    eng = ours.vit_engine(dev)
    eng.profile_begin()
    ours(x)
    ms, cnt, rows = eng.profile_end()
    print(f"GEMM launches {cnt}, {ms:.3f} ms of the step")
    agg = {}
    for key, cin, nt, ks, launches, t in rows:
        kind = key.split(".")[-1] if "layers" in key else "patch_embedding"
        a = agg.setdefault(kind, [0, 0.0, cin, nt])
        a[0] += launches; a[1] += t
    for kind, (launches, t, cin, nt) in agg.items():
        if launches:
            per_layer = 1 if kind == "patch_embedding" else 12                       # launches per pass over all n images
            fl = 2.0 * cin * nt * n * (49 if kind == "patch_embedding" else 50) * per_layer
            print(f"  {kind:16s} [{cin:4d} -> {nt:4d}] x{launches:3d}: {t:7.3f} ms  {fl / t / 1e9:7.1f} TFLOP/s")


if __name__ == "__main__":
    main()
