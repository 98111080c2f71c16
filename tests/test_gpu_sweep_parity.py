"""BASELINE.json config 1 on the GPU (16 synthetic 256x256 RGB images, 3 classes) through BOTH reference drivers, every
level, against the fp32 oracle fed the SAME replayed random draws:

  * Driver B — scripts/robustness_evaluation.py:105-127: all 80 (family, level) pairs through
    `run_sweep(parity_rng=True)`, batch 8 (two full batches);
  * Driver A — models/model_wrappers.py:740-764: all 79 levels through `TestWrapper.test_robustness()`, batch 10
    (one full batch and a ragged one of 6).

Per level: the perturbed network input is bit-equal to the oracle's; the confusion counts are exactly the counts of the
library's own argmax map and differ from the oracle's by no more than the pixels whose argmax differs; mean IoU /
PixelAccuracy / 2I/(1+I) / soft Dice are within 0.1 point (1e-3) of the oracle's; argmax agreement is reported per level
and gated (>= 99.9 % pooled over the sweep).

The oracle network runs on the GPU in IEEE fp32 here (TF32 off, tests/conftest.py) because 159 levels x 16 images on
the CPU would take minutes; `test_gpu_ieee_oracle_equals_cpu_oracle` ties that to the CPU fp32 oracle.
"""
import csv
import os
import random

import numpy as np
import pytest
import torch

from oracle import metrics_oracle as mo
from oracle import perturb_oracle as po
from oracle import unet_oracle
from oracle.make_golden import WRAPPER_LEVELS
from tests import util

pytestmark = pytest.mark.gpu

N, HW = 16, 256
POINT1 = 1e-3  # "0.1 point" on the 0..1 scale the reference reports


@pytest.fixture(scope="module")
def data():
    from image_segmentation_b200 import synthetic
    imgs, masks = synthetic.make_dataset(N, HW, HW, seed=101)
    return imgs, masks


def per_batch_scores(logits, masks, ref_batch):
    """The reference's per-batch metrics and their Python-float means (model_wrappers.py:456-468,
    robustness_evaluation.py:121-124) from oracle logits; also the oracle's argmax map and per-batch counts."""
    n = logits.shape[0]
    pred = unet_oracle.hard_prediction(logits)
    ious, pas, wds, dcs, cms = [], [], [], [], []
    for b in range(-(-n // ref_batch)):
        sl = slice(b * ref_batch, min((b + 1) * ref_batch, n))
        cm = mo.confusion_matrix(pred[sl].numpy(), masks[sl].numpy())
        iou = mo.iou_from_cm(cm)
        ious.append(float(iou)); pas.append(float(mo.pixel_accuracy_from_cm(cm)))
        wds.append(float(mo.wrapper_dice_from_iou(iou)))
        dcs.append(float(np.float32(mo.dice_smp(logits[sl], masks[sl].long()))))
        cms.append(cm)
    mean = lambda v: sum(v) / len(v)
    return dict(iou=mean(ious), pa=mean(pas), wdice=mean(wds), dice=mean(dcs), pred=pred, cm=np.stack(cms))


def check_counts(cm_gpu, pred_gpu, masks, ref_batch, oracle):
    """Counts == counts of the library's own argmax map (bit-exact); distance to the oracle's counts bounded by the
    number of pixels whose argmax differs."""
    n = pred_gpu.shape[0]
    cm_gpu = cm_gpu.cpu().numpy()
    differ = 0
    for b in range(-(-n // ref_batch)):
        sl = slice(b * ref_batch, min((b + 1) * ref_batch, n))
        assert np.array_equal(cm_gpu[b], mo.confusion_matrix(pred_gpu[sl].numpy(), masks[sl].numpy()))
        d = int((pred_gpu[sl].long() != oracle["pred"][sl]).sum())
        assert int(np.abs(cm_gpu[b] - oracle["cm"][b]).sum()) <= 2 * d
        differ += d
    assert int(cm_gpu.sum()) == n * HW * HW
    return differ


def test_gpu_ieee_oracle_equals_cpu_oracle(trained, dev, data):
    """The sweep tests evaluate the oracle on the GPU with TF32 off; tie that to the CPU fp32 oracle on config 1's input."""
    from image_segmentation_b200 import synthetic
    m, sd = trained
    x = synthetic.to_float_nchw(data[0])
    cpu = util.oracle_logits(sd, x)
    gpu = util.oracle_logits_gpu_ieee(sd, x, dev)
    rel = util.rel_err(gpu, cpu)
    agree = util.argmax_agreement(gpu, cpu)
    print(f"oracle on the GPU (IEEE fp32, TF32 off) vs on the CPU: rel err {rel:.3e}, argmax agreement {agree:.6f}")
    assert rel < 2e-5 and agree > 0.9999
    got = m(x.to(dev)).cpu()
    print(f"config 1 (16 x 256x256, no perturbation): library vs CPU fp32 oracle rel err {util.rel_err(got, cpu):.3e}, "
          f"argmax agreement {util.argmax_agreement(got, cpu):.5f}")
    assert util.rel_err(got, cpu) < 1e-2 and util.argmax_agreement(got, cpu) >= 0.999


def _oracle_u8_level(fam, level, u8):
    """Replays the reference wrappers' host draws for one pass over the dataset (customDatasets/perturbations.py) and
    applies the oracle's integer maps.  Consumes `random` / torch's CPU generator exactly like the reference."""
    n, _, h, w = u8.shape
    if fam == "gaussian_noise":
        return np.stack([po.u8_gaussian_noise(u8[i], torch.normal(mean=0.0, std=float(level), size=(3, h, w),
                                                                  dtype=torch.float).numpy()) for i in range(n)])
    if fam == "gaussian_blur":
        return po.u8_blur(u8, level)
    if fam in ("contrast_increase", "contrast_decrease"):
        return po.u8_contrast(u8, level)
    if fam == "brightness_increase":
        return po.u8_brightness(u8, level, True)
    if fam == "brightness_decrease":
        return po.u8_brightness(u8, level, False)
    if fam == "occlusion":
        return np.stack([po.u8_occlusion(u8[i], level, *po.replay_occlusion_corner(level, h, w)) for i in range(n)])
    if fam == "salt_pepper_noise":
        return np.stack([po.u8_salt_pepper(u8[i], po.replay_salt_pepper_code(level, h, w)) for i in range(n)])
    raise KeyError(fam)


def test_script_sweep_all_80_levels_config1(trained, dev, data, tmp_path):
    from image_segmentation_b200.scripts import robustness_evaluation as R
    m, sd = trained
    imgs, masks = data
    out = str(tmp_path / "results" / "robustness_scores.csv")
    rows = R.run_sweep(m, imgs, masks, results_file=out, batch_size=8, parity_rng=True, quiet=True, device=dev)
    with open(out) as f:
        lines = list(csv.reader(f))
    assert lines[0] == ["perturbation_type", "param_value", "mean_dice"] and len(lines) == 81
    got_rows = {(r[0], i % 10): r[2] for i, r in enumerate(rows)}

    # second pass: same seeds (robustness_evaluation.py:32-34), oracle and library fed the same draws level by level
    random.seed(42); np.random.seed(42); torch.manual_seed(42)
    eng = m.engine(dev)
    di, dm = imgs.to(dev), masks.to(dev)
    u8 = imgs.permute(0, 3, 1, 2).numpy()
    differ = total = 0
    worst, gaps = [], []
    for fam, levels in R.PERTURBATIONS.items():
        for li, level in enumerate(levels):
            state = (random.getstate(), torch.get_rng_state())
            want_u8 = _oracle_u8_level(fam, level, u8)
            random.setstate(state[0]); torch.set_rng_state(state[1])
            spec = R.parity_spec(fam, level, N, HW, HW)
            res = eng.eval_batches(di, dm, spec, 8, want_dice=True, want_pred=True, want_perturbed=True)
            want_x = po.dequantise(want_u8)
            assert np.array_equal(res["perturbed"].cpu().numpy(), want_x), (fam, level)   # bit-equal network input
            o = per_batch_scores(util.oracle_logits_gpu_ieee(sd, torch.from_numpy(want_x), dev), masks, 8)
            pred = res["pred"].cpu()
            d = check_counts(res["cm"], pred, masks, 8, o)
            differ += d; total += pred.numel()
            worst.append((1.0 - d / pred.numel(), fam, level))
            from image_segmentation_b200.evaluator import RobustnessEvaluator as E
            iou, pa, wd = E.wrapper_scores(res["cm"])
            dice = E.script_score(res["dice"])
            assert abs(dice - got_rows[(fam, li)]) < 1e-5, (fam, level)                    # what the CSV holds
            for name, a, b in (("IoU", iou, o["iou"]), ("PixelAccuracy", pa, o["pa"]), ("2I/(1+I)", wd, o["wdice"]),
                               ("soft Dice", dice, o["dice"])):
                gaps.append((abs(a - b), fam, level, name))
    worst.sort()
    gaps.sort(reverse=True)
    print(f"script sweep, 80 levels x {N} images: pooled argmax agreement {1 - differ / total:.6f}; lowest levels: "
          + ", ".join(f"{f} {l}: {a:.5f}" for a, f, l in worst[:4]) + "; largest metric gaps to the oracle: "
          + ", ".join(f"{f} {l} {n}: {d:.2e}" for d, f, l, n in gaps[:3]))
    assert gaps[0][0] < POINT1, gaps[:3]
    assert 1 - differ / total >= 0.999
    assert worst[0][0] >= 0.995


WRAPPER_FILES = [  # (csv file, class name, slice of WRAPPER_LEVELS[class]) in test_robustness order
    ("gaussian_pixel_noise_results.csv", "GaussianPixelNoise", slice(0, 10)),
    ("repeated_blur_results.csv", "RepeatedBlur", slice(0, 10)),
    ("contrast_increase_results.csv", "ContrastChange", slice(0, 10)),
    ("contrast_decrease_results.csv", "ContrastChange", slice(10, 20)),
    ("brightness_increase_results.csv", "BrightnessChange", slice(0, 20)),  # the reference passes the negative offsets
    ("occlusion_results.csv", "Occlusion", slice(0, 10)),                    # with increase=True: one file, 20 rows
    ("salt_and_pepper_results.csv", "SaltAndPepper", slice(0, 9)),
]


def _oracle_f_batch(cls_name, level, x, aux):
    """One batch through the oracle's float-space maps (models/processing_blocks.py:471-474,493-496,516-518,535-539,
    557-563,586-592), fed the numbers the drop-in module drew for this batch (`aux`: randn / corners / rand)."""
    if cls_name == "GaussianPixelNoise":
        return po.f_gaussian_noise(x, aux.cpu().numpy(), level)
    if cls_name == "RepeatedBlur":
        return po.f_box_blur(x, level)
    if cls_name == "ContrastChange":
        return po.f_contrast(x, level)
    if cls_name == "BrightnessChange":
        return po.f_brightness(x, level)
    if cls_name == "Occlusion":
        return po.f_occlusion(x, int(level), [tuple(int(v) for v in c) for c in aux.cpu().numpy()])
    if cls_name == "SaltAndPepper":
        return po.f_salt_pepper(x, aux.cpu().numpy(), level)
    raise KeyError(cls_name)


def test_testwrapper_robustness_all_79_levels_config1(trained, dev, data, tmp_path, monkeypatch):
    from image_segmentation_b200 import runtime, synthetic
    from image_segmentation_b200.customDatasets.datasets import CustomImageDataset
    from image_segmentation_b200.models import processing_blocks as PB
    from image_segmentation_b200.models.model_wrappers import TestWrapper
    from image_segmentation_b200.models.UNet import UNet
    m, sd = trained
    imgs, masks = data
    monkeypatch.chdir(tmp_path)
    os.makedirs("augmentation-results")  # the reference requires the directory to exist (model_wrappers.py:502-506)
    path = str(tmp_path / "model.pth")
    torch.save({"_orig_mod." + k: v for k, v in sd.items()}, path)
    tw = TestWrapper(UNet(), test_dataset_class=CustomImageDataset,
                     test_dataset_args={"split": "test", "augmentations_per_datapoint": 0, "cache": False, "n": N,
                                        "h": HW, "w": HW, "seed": 101}, batch_size=10, model_load_location=path)
    tw.evaluator()                      # (the loader that makes the set resident draws its base seed: before seeding)
    random.seed(7); torch.manual_seed(7)
    tw.test_robustness()
    levels_all = dict(WRAPPER_LEVELS, RepeatedBlur=[0, 1, 2, 3, 4, 5, 6, 7, 8, 9])
    csv_rows = {}
    for fname, cls_name, sl in WRAPPER_FILES:
        with open(os.path.join("augmentation-results", fname)) as f:
            rows = list(csv.DictReader(f))
        assert len(rows) == len(levels_all[cls_name][sl]), fname
        csv_rows[(fname, cls_name)] = rows
    assert sum(len(r) for r in csv_rows.values()) == 79

    # second pass with the same seeds: the drop-in modules draw, level by level and batch by batch, the numbers they drew
    # inside test_robustness (same generators, same order — where the reference draws them); the oracle is fed exactly
    # those numbers, so inputs can be compared bit for bit and the CSV rows against the oracle's metrics
    random.seed(7); torch.manual_seed(7)
    x_all = synthetic.to_float_nchw(imgs)
    eng = tw.model.engine(dev)
    dm = masks.to(dev)
    differ = total = 0
    worst, gaps = [], []
    for fname, cls_name, sl in WRAPPER_FILES:
        for row, level in zip(csv_rows[(fname, cls_name)], levels_all[cls_name][sl]):
            assert float(row["Parameter Value"]) == float(level)
            want = []
            got = []
            for b0 in range(0, N, 10):
                xb = x_all[b0:b0 + 10]
                mod = getattr(PB, cls_name)(level)
                xd = xb.to(dev).contiguous()   # (to_float_nchw keeps the NHWC strides of its uint8 source)
                spec = mod.spec(xd)                                   # the module's own draws for this batch
                got.append(runtime.engine_for(dev).perturb(xd, spec).cpu().numpy())   # = what mod.forward computes
                want.append(_oracle_f_batch(cls_name, level, xb.contiguous().numpy(), spec.aux))
            want, got = np.concatenate(want), np.concatenate(got)
            assert np.array_equal(got, want), (cls_name, level, int((got != want).sum()),  # bit-equal network input
                                               float(np.abs(got - want).max()))
            o = per_batch_scores(util.oracle_logits_gpu_ieee(sd, torch.from_numpy(want), dev), masks, 10)
            res = eng.eval_batches(torch.from_numpy(want).to(dev), dm, None, 10, want_pred=True)
            pred = res["pred"].cpu()
            d = check_counts(res["cm"], pred, masks, 10, o)
            differ += d; total += pred.numel()
            worst.append((1.0 - d / pred.numel(), cls_name, level))
            for col, key in (("Avg IoU Loss", "iou"), ("Avg Pixel Accuracy Loss", "pa"), ("Avg Dice Loss", "wdice")):
                gaps.append((abs(float(row[col]) - o[key]), cls_name, level, col))
    worst.sort()
    gaps.sort(reverse=True)
    print(f"TestWrapper.test_robustness, 79 levels x {N} images: pooled argmax agreement {1 - differ / total:.6f}; "
          "lowest levels: " + ", ".join(f"{f} {l}: {a:.5f}" for a, f, l in worst[:4]) + "; largest metric gaps to the oracle: "
          + ", ".join(f"{f} {l} {n}: {d:.2e}" for d, f, l, n in gaps[:3]))
    assert gaps[0][0] < POINT1, gaps[:3]
    assert 1 - differ / total >= 0.999
    assert worst[0][0] >= 0.995
