"""The two reference drivers end to end on the GPU (small synthetic set): CSV schema, per-batch metric
semantics, and values against the oracle run on the same inputs (within 0.1 point)."""
import csv
import os

import numpy as np
import pytest
import torch

from oracle import metrics_oracle as mo
from oracle import perturb_oracle as po
from oracle import unet_oracle
from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup(tmp_path_factory):
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.UNet import UNet
    torch.manual_seed(42)
    m = UNet().eval()
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    imgs, masks = synthetic.make_dataset(32, 128, 128, seed=21)
    sd = util.train_like(sd, synthetic.to_float_nchw(imgs), masks, steps=200, lr=1e-3, batch=8, crop=128)
    d = tmp_path_factory.mktemp("ckpt")
    path = str(d / "model.pth")
    torch.save({"_orig_mod." + k: v for k, v in sd.items()}, path)  # torch.compile prefix, as the reference saves
    return sd, path, d


def oracle_level(sd, imgs_u8_nhwc, masks, pert_fn, ref_batch, dev):
    x_u8 = imgs_u8_nhwc.permute(0, 3, 1, 2).numpy()
    x = torch.from_numpy(po.dequantise(pert_fn(x_u8)))
    logits = unet_oracle.forward({k: v.to(dev) for k, v in sd.items()}, x.to(dev)).cpu()
    n = x.shape[0]
    ious, pas, dices = [], [], []
    for b in range(-(-n // ref_batch)):
        sl = slice(b * ref_batch, min((b + 1) * ref_batch, n))
        cm = mo.confusion_matrix(unet_oracle.hard_prediction(logits[sl]).numpy(), masks[sl].numpy())
        ious.append(float(mo.iou_from_cm(cm)))
        pas.append(float(mo.pixel_accuracy_from_cm(cm)))
        dices.append(mo.dice_smp(logits[sl], masks[sl].long()))
    return np.mean(ious), np.mean(pas), np.mean(dices)


def test_script_sweep_csv(setup):
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.UNet import UNet
    from image_segmentation_b200.scripts import robustness_evaluation as R
    sd, path, d = setup
    dev = torch.device("cuda:0")
    m = UNet().eval()
    m.load_state_dict(sd)
    m = m.to(dev)
    imgs, masks = synthetic.make_dataset(21, 64, 128, seed=31)  # ragged last batch of 5
    out = str(d / "results" / "robustness_scores.csv")
    subset = {"gaussian_blur": [0, 3], "contrast_decrease": [1.0, 0.4], "brightness_increase": [0, 45],
              "salt_pepper_noise": [0.0, 0.06], "occlusion": [0, 25], "gaussian_noise": [0, 10]}
    rows = R.run_sweep(m, imgs, masks, results_file=out, batch_size=8, perturbations=subset, quiet=True)
    with open(out) as f:
        lines = list(csv.reader(f))
    assert lines[0] == ["perturbation_type", "param_value", "mean_dice"]
    assert len(lines) == 1 + 12 and all(len(r) == 3 for r in lines)
    float(lines[1][1]); float(lines[1][2])  # what scripts/plot_robustness_evaluation.py:36-48 parses
    got = {(r[0], str(r[1])): r[2] for r in rows}
    checks = {("gaussian_blur", "3"): lambda u: po.u8_blur(u, 3), ("contrast_decrease", "0.4"): lambda u: po.u8_contrast(u, 0.4),
              ("brightness_increase", "45"): lambda u: po.u8_brightness(u, 45, True), ("gaussian_blur", "0"): lambda u: u}
    for key, fn in checks.items():
        _, _, dice = oracle_level(sd, imgs, masks, fn, 8, dev)
        assert abs(got[key] - dice) < 1e-3, (key, got[key], dice)  # 0.1 point
    assert got[("gaussian_noise", "0")] == got[("gaussian_blur", "0")] == got[("occlusion", "0")]
    assert got[("salt_pepper_noise", "0.06")] < got[("salt_pepper_noise", "0.0")] + 1e-3


def test_testwrapper_csvs(setup, monkeypatch):
    from image_segmentation_b200.customDatasets.datasets import CustomImageDataset
    from image_segmentation_b200.models.model_wrappers import TestWrapper
    from image_segmentation_b200.models.UNet import UNet
    from image_segmentation_b200 import synthetic
    sd, path, d = setup
    monkeypatch.chdir(d)
    dev = torch.device("cuda:0")
    with pytest.raises(AssertionError):
        TestWrapper(UNet())
    tw = TestWrapper(UNet(), test_dataset_class=CustomImageDataset,
                     test_dataset_args={"split": "test", "augmentations_per_datapoint": 0, "cache": False, "n": 23,
                                        "h": 64, "w": 128, "seed": 33}, batch_size=10, model_load_location=path)
    tw.test()
    tw.test_contrast_change([1.0, 0.4], False)
    tw.test_brightness_change([0, -30], True)
    tw.test_occlusion([0, 20])
    tw.test_gaussian_pixel_noise([1e-6, 10])
    tw.test_salt_and_pepper([0.0, 0.1])
    tw.test_repeated_blur([0, 2])
    files = sorted(os.listdir(d / "augmentation-results"))
    assert files == ["brightness_increase_results.csv", "contrast_decrease_results.csv", "gaussian_pixel_noise_results.csv",
                     "occlusion_results.csv", "repeated_blur_results.csv", "salt_and_pepper_results.csv"]
    rows = list(csv.DictReader(open(d / "augmentation-results" / "contrast_decrease_results.csv")))
    assert list(rows[0].keys()) == ["Augmentation", "Parameter Value", "Avg IoU Loss", "Avg Pixel Accuracy Loss", "Avg Dice Loss"]
    assert rows[0]["Augmentation"] == "Contrast Change" and rows[1]["Parameter Value"] == "0.4"
    imgs, masks = synthetic.make_dataset(23, 64, 128, seed=33)
    x = synthetic.to_float_nchw(imgs).numpy()
    for row, fn in ((rows[0], lambda a: a), (rows[1], lambda a: po.f_contrast(a, 0.4))):
        logits = unet_oracle.forward({k: v.to(dev) for k, v in sd.items()}, torch.from_numpy(fn(x)).to(dev)).cpu()
        ious, pas = [], []
        for b in range(3):
            sl = slice(b * 10, min((b + 1) * 10, 23))
            cm = mo.confusion_matrix(unet_oracle.hard_prediction(logits[sl]).numpy(), masks[sl].numpy())
            ious.append(float(mo.iou_from_cm(cm))); pas.append(float(mo.pixel_accuracy_from_cm(cm)))
        assert abs(float(row["Avg IoU Loss"]) - np.mean(ious)) < 1e-3
        assert abs(float(row["Avg Pixel Accuracy Loss"]) - np.mean(pas)) < 1e-3
    tw.test_contrast_change([1.0], False)  # append mode: header written once
    rows2 = list(csv.reader(open(d / "augmentation-results" / "contrast_decrease_results.csv")))
    assert len(rows2) == 4 and rows2[0][0] == "Augmentation"


def test_testwrapper_on_device_resident_rows(setup, monkeypatch):
    """TestWrapper fed by DeviceResidentDataset (rows decoded on the GPU) gives the confusion counts of the generic
    loader path on the same decoded data."""
    from image_segmentation_b200.customDatasets.datasets import DeviceResidentDataset
    from image_segmentation_b200.models.model_wrappers import TestWrapper
    from image_segmentation_b200.models.UNet import UNet
    from image_segmentation_b200 import synthetic
    from oracle import dataset_oracle as do
    sd, path, d = setup
    monkeypatch.chdir(d)
    rows = synthetic.make_raw_rows(7, 256, 256, seed=4)

    class HostRows(torch.utils.data.Dataset):          # what the reference's CPU __getitem__ yields
        def __init__(self):
            self.items = [do.deserialize_datapoint(r) for r in rows]

        def __len__(self):
            return len(self.items)

        def __getitem__(self, i):
            return torch.from_numpy(self.items[i][0]), torch.from_numpy(self.items[i][1])

    a = TestWrapper(UNet(), test_dataset_class=lambda: DeviceResidentDataset.from_rows(rows, "cuda:0"),
                    test_dataset_args={}, batch_size=4, model_load_location=path)
    b = TestWrapper(UNet(), test_dataset_class=HostRows, test_dataset_args={}, batch_size=4, model_load_location=path)
    cm_a = a.evaluator().run_level(None)[0].clone()
    cm_b = b.evaluator().run_level(None)[0].clone()
    assert cm_a.shape == (2, 3, 3) and torch.equal(cm_a, cm_b) and int(cm_a.sum()) == 7 * 256 * 256
