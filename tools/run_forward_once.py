"""One warm-up and one timed evaluation level over N synthetic 256x256 images (default 148 = one micro-batch): the
process ncu attaches to when a single kernel of the U-Net path is captured.  Usage: python tools/run_forward_once.py [N]"""
import sys

import torch

sys.path.insert(0, ".")


def main():
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.evaluator import RobustnessEvaluator
    from image_segmentation_b200.models.UNet import UNet
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 148
    dev = torch.device("cuda:0")
    torch.manual_seed(42)
    m = UNet().eval().to(dev)
    imgs, masks = synthetic.make_dataset(16, 256, 256, seed=3)
    reps = (n + 15) // 16
    ev = RobustnessEvaluator(m, imgs.repeat(reps, 1, 1, 1)[:n], masks.repeat(reps, 1, 1)[:n], 8, dev, rank=0, world_size=1)
    for _ in range(2):
        cm, _ = ev.run_level(None)
    torch.cuda.synchronize()
    print("pixels counted", int(cm.sum()))


if __name__ == "__main__":
    main()
