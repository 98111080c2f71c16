"""The reference's customDatasets/datasets.py reads the Oxford-IIIT-Pet rows through HuggingFace
`datasets`, which needs the network; only its OUTPUT CONTRACT (:137-160) is on this path, and
`SyntheticSegmentationDataset` honours it.  `CustomImageDataset` keeps the reference's name and
argument checks so calling code runs unchanged on synthetic data."""
from ..synthetic import SyntheticSegmentationDataset, make_dataset  # noqa: F401


class CustomImageDataset(SyntheticSegmentationDataset):
    def __init__(self, split="train", augmentations_per_datapoint=0, cache=False, n=64, h=256, w=256, seed=99, **kw):
        if split not in ("train", "validation", "test"):  # datasets.py:33-34
            raise ValueError(f"Invalid split: {split}. Must be one of 'train', 'validation', or 'test'.")
        super().__init__(n=n, h=h, w=w, seed=seed)
