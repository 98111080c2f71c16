"""The reference's customDatasets/datasets.py reads the Oxford-IIIT-Pet rows through HuggingFace
`datasets`, which needs the network; only its OUTPUT CONTRACT (:137-160) is on this path, and
`SyntheticSegmentationDataset` honours it.  `CustomImageDataset` keeps the reference's name and
argument checks so calling code runs unchanged on synthetic data."""
import numpy as np
import torch

from ..engine import Perturbation

from ..synthetic import SyntheticSegmentationDataset, make_dataset  # noqa: F401


class CustomImageDataset(SyntheticSegmentationDataset):
    def __init__(self, split="train", augmentations_per_datapoint=0, cache=False, n=64, h=256, w=256, seed=99, **kw):
        if split not in ("train", "validation", "test"):  # datasets.py:33-34
            raise ValueError(f"Invalid split: {split}. Must be one of 'train', 'validation', or 'test'.")
        super().__init__(n=n, h=h, w=w, seed=seed)


class DeviceResidentDataset(torch.utils.data.Dataset):
    """The reference `CustomImageDataset`'s rows (customDatasets/datasets.py:92-160) decoded ON THE DEVICE and
    kept there in the evaluation form: images uint8 [n,H,W,3] (the serialized bytes as they are), masks uint8
    [n,H,W] in {0,1,2} (the trimap remap of `_deserialize_datapoint`, :118-131, done by `ub_decode_masks`).

    * `from_rows(rows, device)`  — rows with the HuggingFace schema the reference reads (`row['image']`: H*W*3 bytes,
      `row['mask']`: H*W bytes; :107-110): bytes are concatenated into one pinned buffer, uploaded once, decoded by
      one kernel launch.  No per-sample CPU work, no per-batch H2D.
    * `from_cache(items, device)` — the reference's `{split}_dataset.pt` cache (:64-86): a list of
      (fp32 [3,H,W] = k/255, int64 [H,W]) tuples, or the path of such a file; uploaded in chunks and packed to
      uint8 by `ub_pack_dataset`.
    * `resident_tensors()` -> (images_u8, masks_u8): what `evaluator.RobustnessEvaluator` takes.
    * `__len__` / `__getitem__(idx)` keep the reference's item contract — (fp32 [3,H,W] in [0,1], int64 [H,W]),
      `idx // (augmentations_per_datapoint + 1)` selecting the row (:150-158) — on the device.
    """

    def __init__(self, images_u8, masks_u8, augmentations_per_datapoint=0):
        assert isinstance(augmentations_per_datapoint, int) and augmentations_per_datapoint >= 0, \
            f"You must choose a positive integer for augmentations per datapoint, you choose: {augmentations_per_datapoint}"
        if not images_u8.is_cuda:
            raise RuntimeError("DeviceResidentDataset holds CUDA tensors (unet-b200 has no CPU path)")
        self.images_u8, self.masks_u8 = images_u8, masks_u8
        self.augmentations_per_datapoint = augmentations_per_datapoint + 1

    @classmethod
    def from_rows(cls, rows, device, h=256, w=256, augmentations_per_datapoint=0):
        from .. import runtime
        rows = list(rows)
        n = len(rows)
        img_host = torch.empty((n, h, w, 3), dtype=torch.uint8).pin_memory()
        raw_host = torch.empty((n, h, w), dtype=torch.uint8).pin_memory()
        iv, mv = img_host.view(n, -1).numpy(), raw_host.view(n, -1).numpy()
        for i, row in enumerate(rows):
            ib, mb = row["image"], row["mask"]
            if len(ib) != h * w * 3 or len(mb) != h * w:
                raise ValueError(f"row {i}: expected {h * w * 3} image bytes and {h * w} mask bytes, "
                                 f"got {len(ib)} and {len(mb)}")
            iv[i] = np.frombuffer(ib, dtype=np.uint8)
            mv[i] = np.frombuffer(mb, dtype=np.uint8)
        dev = torch.device(device)
        images = img_host.to(dev, non_blocking=True)
        masks = runtime.engine_for(dev).decode_masks(raw_host.to(dev, non_blocking=True))
        return cls(images, masks, augmentations_per_datapoint)

    @classmethod
    def from_cache(cls, items, device, chunk=256, augmentations_per_datapoint=0):
        from .. import runtime
        if isinstance(items, (str, bytes)) or hasattr(items, "__fspath__"):
            items = torch.load(items, weights_only=True)  # datasets.py:72
        dev = torch.device(device)
        eng = runtime.engine_for(dev)
        imgs, masks = [], []
        for i0 in range(0, len(items), chunk):
            part = items[i0:i0 + chunk]
            x = torch.stack([p[0] for p in part]).float()
            t = torch.stack([torch.as_tensor(p[1]) for p in part]).long()
            a, b = eng.pack_dataset(x.to(dev), t.to(dev))
            back = eng.perturb(a, Perturbation())  # u8 -> fp32 with the reference's true division by 255
            if not torch.equal(back, x.to(dev)):
                raise RuntimeError("cached images are not k/255: the device-resident uint8 form does not apply")
            imgs.append(a)
            masks.append(b)
        return cls(torch.cat(imgs), torch.cat(masks), augmentations_per_datapoint)

    def resident_tensors(self):
        return self.images_u8, self.masks_u8

    def __len__(self):
        return self.images_u8.shape[0] * self.augmentations_per_datapoint

    def __getitem__(self, idx):
        from .. import runtime
        i = idx // self.augmentations_per_datapoint
        # u8/255 in the library (IEEE division like the CPU reference; torch's CUDA `/ 255.0` multiplies by 1/255)
        x = runtime.engine_for(self.images_u8.device).perturb(self.images_u8[i:i + 1].contiguous(), Perturbation())
        return x[0], self.masks_u8[i].long()
