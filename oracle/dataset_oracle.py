"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of the reference's row decoding,
customDatasets/datasets.py:92-134 (`CustomImageDataset._deserialize_datapoint` / `_deserialize_numpy`).

Pinned against the reference's own method (run unbound on synthetic rows by oracle/make_golden.py ->
tests/golden/dataset_rows.npz)."""
import numpy as np


def deserialize_numpy(byte_data, shape, dtype=np.uint8):
    """datasets.py:133-134: reinterpret the serialized bytes."""
    return np.frombuffer(byte_data, dtype=dtype).reshape(shape).copy()


def decode_mask(raw: np.ndarray) -> np.ndarray:
    """datasets.py:118-131.  raw uint8 [H,W] trimap -> int64 class ids.  Per image: if any pixel is 38 (cat),
    1 where raw is 38 or 255; otherwise 2 where raw is 75 or 255; 0 elsewhere."""
    cat = (raw == 38).astype(np.int64)
    dog = np.where(raw == 75, 2, 0).astype(np.int64)
    unc = (raw == 255).astype(np.int64)
    return cat + unc if cat.sum() > 0 else dog + 2 * unc


def deserialize_datapoint(row, h=256, w=256):
    """datasets.py:107-131 -> (image fp32 [3,H,W] = u8/255, mask int64 [H,W])."""
    image = deserialize_numpy(row["image"], (h, w, 3))
    raw = deserialize_numpy(row["mask"], (h, w))
    return image.transpose(2, 0, 1).astype(np.float32) / np.float32(255.0), decode_mask(raw)
