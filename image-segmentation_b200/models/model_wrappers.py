"""Drop-in for `TestWrapper` of the reference's models/model_wrappers.py:251-792 (the evaluation
driver; the training wrappers are out of scope).  Same constructor, same public methods, same CSV files
(`augmentation-results/<family>_results.csv`, fields and append mode of :502-521), same per-batch
metric semantics (:456-468) — the loop body runs as one fused call per chunk of batches on the GPU."""
from __future__ import annotations

import csv
import os
from collections import OrderedDict

import torch
from torch.utils.data import DataLoader
from tqdm import tqdm

from ..customDatasets.datasets import CustomImageDataset
from ..engine import Perturbation
from ..evaluator import RobustnessEvaluator
from .processing_blocks import (BrightnessChange, ContrastChange, GaussianPixelNoise, Occlusion, RepeatedBlur,
                                SaltAndPepper, _DevicePerturbation)


class TestWrapper:
    __test__ = False  # not a pytest class

    def __init__(self, model, test_dataset_class=CustomImageDataset,
                 test_dataset_args={'split': 'test', 'augmentations_per_datapoint': 0, 'cache': False},
                 batch_size=10, model_load_location=None):
        assert model_load_location is not None, "Model load location must be specified"
        self.batch_size = batch_size
        self.test_dataset = test_dataset_class(**test_dataset_args)
        self.test_dataloader = DataLoader(self.test_dataset, batch_size=batch_size, shuffle=False, pin_memory=False)
        state_dict = torch.load(model_load_location, map_location="cpu", weights_only=True)
        new_state_dict = OrderedDict((k.replace("_orig_mod.", ""), v) for k, v in state_dict.items())
        model.load_state_dict(new_state_dict)
        if not torch.cuda.is_available():
            raise RuntimeError("unet-b200 needs a CUDA sm_100a device (no CPU fallback)")
        self.device = torch.device('cuda', torch.cuda.current_device())
        self.model = model.to(self.device).eval()
        self._evaluator = None

    # ---- dataset made device-resident once (uint8 when the items are exact k/255, as the reference's are)
    def evaluator(self) -> RobustnessEvaluator:
        if self._evaluator is None and hasattr(self.test_dataset, "resident_tensors"):
            # customDatasets.datasets.DeviceResidentDataset: rows were decoded on the device already
            u8, t = self.test_dataset.resident_tensors()
            self._evaluator = RobustnessEvaluator(self.model, u8, t, self.batch_size, self.device)
        if self._evaluator is None:
            imgs, masks = [], []
            for x, t in DataLoader(self.test_dataset, batch_size=64, shuffle=False):
                imgs.append(x)
                masks.append(t)
            x = torch.cat(imgs).float()
            t = torch.cat(masks)
            u8 = (x * 255.0).round().clamp(0, 255).to(torch.uint8)
            if not torch.equal(u8.float() / 255.0, x):
                raise RuntimeError("dataset images are not k/255: the device-resident uint8 path does not apply")
            self._evaluator = RobustnessEvaluator(self.model, u8.permute(0, 2, 3, 1).contiguous(), t.to(torch.uint8),
                                                  self.batch_size, self.device)
        return self._evaluator

    def _report(self, cm, header=None):
        iou, pa, dc = RobustnessEvaluator.wrapper_scores(cm)
        if header:
            tqdm.write(header)
        tqdm.write(f"IoU: {iou:.4f}")
        tqdm.write(f"Pixel Accuracy: {pa:.4f}")
        tqdm.write(f"Dice: {dc:.4f}")
        tqdm.write('\n')
        return iou, pa, dc

    def test(self):
        cm, _ = self.evaluator().run_level(None)
        self._report(cm)

    def test_augmentation(self, augmentation, param_val):
        ev = self.evaluator()
        self.model.eval()
        if isinstance(augmentation, _DevicePerturbation):
            cm, _ = ev.run_level_per_batch(augmentation.spec)
        else:
            # foreign nn.Module: apply it batch by batch in torch, then the fused network+metrics call
            augmentation.to(self.device)
            ev.cm.zero_()
            with torch.no_grad():
                for b in range(ev.b0, ev.b1):
                    lo, hi = b * ev.ref_batch - ev.i0, min((b + 1) * ev.ref_batch, ev.n_total) - ev.i0
                    x = ev.images[lo:hi].permute(0, 3, 1, 2).float() / 255.0
                    x = augmentation(x).contiguous()
                    ev.engine.eval_batches(x, ev.masks[lo:hi], None, ev.ref_batch, out_cm=ev.cm[b:b + 1])
            cm = ev.cm
        return self._report(cm, f"Augmentation: {augmentation.__class__.__name__}, Parameter Value: {param_val}")

    def log_results_to_csv(self, results, filename=None):
        os.makedirs('augmentation-results', exist_ok=True)
        filename = 'augmentation-results/' + filename
        fieldnames = ['Augmentation', 'Parameter Value', 'Avg IoU Loss', 'Avg Pixel Accuracy Loss', 'Avg Dice Loss']
        with open(filename, mode='a', newline='') as file:
            writer = csv.DictWriter(file, fieldnames=fieldnames)
            if file.tell() == 0:
                writer.writeheader()
            for r in results:
                writer.writerow(dict(zip(fieldnames, r)))

    def _family(self, label, make, values, filename):
        results = []
        for v in values:
            results.append((label, v) + tuple(self.test_augmentation(make(v), v)))
        self.log_results_to_csv(results, filename=filename)

    def test_gaussian_pixel_noise(self, parameter_values):
        self._family('Gaussian Pixel Noise', GaussianPixelNoise, parameter_values, 'gaussian_pixel_noise_results.csv')

    def test_repeated_blur(self, parameter_values):
        self._family('Repeated Blur', RepeatedBlur, parameter_values, 'repeated_blur_results.csv')

    def test_contrast_change(self, parameter_values, increase=True):
        self._family('Contrast Change', ContrastChange, parameter_values,
                     'contrast_increase_results.csv' if increase else 'contrast_decrease_results.csv')

    def test_brightness_change(self, parameter_values, increase=True):
        self._family('Brightness Change', BrightnessChange, parameter_values,
                     'brightness_increase_results.csv' if increase else 'brightness_decrease_results.csv')

    def test_occlusion(self, parameter_values):
        self._family('Occlusion', Occlusion, parameter_values, 'occlusion_results.csv')

    def test_salt_and_pepper(self, parameter_values):
        self._family('Salt and Pepper', SaltAndPepper, parameter_values, 'salt_and_pepper_results.csv')

    def test_robustness(self):
        # levels and quirks of model_wrappers.py:740-764 (noise level 0 is 1e-6; brightness decrease is
        # passed as negative offsets with increase=True; salt-and-pepper has 9 levels)
        self.test_gaussian_pixel_noise([1e-6, 2, 4, 6, 8, 10, 12, 14, 16, 18])
        self.test_repeated_blur([0, 1, 2, 3, 4, 5, 6, 7, 8, 9])
        self.test_contrast_change([1.0, 1.01, 1.02, 1.03, 1.04, 1.05, 1.1, 1.15, 1.20, 1.25], True)
        self.test_contrast_change([1.0, 0.95, 0.90, 0.85, 0.80, 0.60, 0.40, 0.30, 0.20, 0.10], False)
        self.test_brightness_change([0, 5, 10, 15, 20, 25, 30, 35, 40, 45], True)
        self.test_brightness_change([0, -5, -10, -15, -20, -25, -30, -35, -40, -45], True)
        self.test_occlusion([0, 5, 10, 15, 20, 25, 30, 35, 40, 45])
        self.test_salt_and_pepper([0.00, 0.02, 0.04, 0.06, 0.08, 0.10, 0.12, 0.14, 0.16])

    def plot_predicitons(self, indicies=None):
        raise NotImplementedError("plotting (matplotlib) is outside the scope of unet-b200")
