import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


import torch

# Checker-side arithmetic must be IEEE fp32: PyTorch lets cuDNN / cuBLAS use TF32 for fp32 convolutions by default,
# whose own error (~1e-3 relative) is of the order of the bf16 tolerance being gated.  Every logits / argmax gate
# evaluates the oracle on the CPU (fp32); the few places that run it on the GPU for speed (the 159-level sweep
# test, the checker-side training) get IEEE fp32 through these switches, and test_gpu_sweep_parity.py checks that the
# GPU-side oracle agrees with the CPU one.
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU")
    config.addinivalue_line("markers", "reference: needs the reference tree at /root/reference (build container only)")


def pytest_collection_modifyitems(config, items):
    from oracle import refload
    if not refload.available():
        skip = pytest.mark.skip(reason="/root/reference not present on this machine")
        for it in items:
            if "reference" in it.keywords:
                it.add_marker(skip)


# ---------------------------------------------------------------- shared GPU fixtures (session scope: one training run)
@pytest.fixture(scope="session")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a B200"
    return torch.device("cuda:0")


@pytest.fixture(scope="session")
def base_model(dev):
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.UNet import UNet
    torch.manual_seed(42)
    m = UNet().eval()
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    m.load_state_dict(sd)
    return m.to(dev), sd


@pytest.fixture(scope="session")
def trained(dev):
    """UNet with decision margins: a few hundred Adam steps on the learnable synthetic task (checker side, IEEE fp32)."""
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.UNet import UNet
    from tests import util
    torch.manual_seed(42)
    m = UNet().eval()
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    imgs, masks = synthetic.make_dataset(48, 128, 128, seed=21)
    sd = util.train_like(sd, synthetic.to_float_nchw(imgs), masks, steps=400, lr=1e-3, batch=8, crop=128, augment=True)
    m.load_state_dict(sd)
    return m.to(dev), sd
