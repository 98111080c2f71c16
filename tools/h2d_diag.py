"""Diagnostic (GPU box): per-step wall times of bench.py's two loops — device-resident levels vs the end-to-end loop
(double-buffered H2D + D2H of the counts) — over the same interleaved (family, level) sequence."""
import sys, os, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from image_segmentation_b200 import synthetic
from image_segmentation_b200.evaluator import RobustnessEvaluator
from image_segmentation_b200.models.UNet import UNet
from image_segmentation_b200.scripts.robustness_evaluation import make_spec
dev = torch.device("cuda:0"); torch.cuda.set_device(dev)
torch.manual_seed(42)
m = UNet().eval().to(dev)
bi, bm = synthetic.make_dataset(230, 256, 256, seed=99)
images = bi.repeat(16, 1, 1, 1)[:3680].contiguous().pin_memory(); masks = bm.repeat(16, 1, 1)[:3680].contiguous().pin_memory()
ev = RobustnessEvaluator(m, images, masks, 8, dev, rank=0, world_size=1)
levels = bench.level_list()
def step(i):
    name, lv = levels[i % len(levels)]
    return ev.run_level(make_spec(name, lv, 65536, seed=i))
for i in range(3): step(i)
torch.cuda.synchronize()
K = 10
for rep in range(2):
    ts = []
    for i in range(K):
        torch.cuda.synchronize(); t0 = time.perf_counter(); step(3 + i); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    print("resident", [levels[(3 + i) % 80][0][:8] + ":" + str(levels[(3 + i) % 80][1]) for i in range(K)])
    print("resident ms", [round(t, 1) for t in ts], "sum", round(sum(ts), 1))
    ev.upload_async(images, masks); torch.cuda.synchronize()
    ts = []
    for i in range(K):
        t0 = time.perf_counter()
        ev.swap()
        if i + 1 < K: ev.upload_async(images, masks)
        cm, _ = step(3 + i); cm.cpu()
        ts.append((time.perf_counter() - t0) * 1e3)
    print("e2e ms     ", [round(t, 1) for t in ts], "sum", round(sum(ts), 1))
