"""Process-wide bits: default micro-batch, and one shared Engine per device for the stand-alone
perturbation / metric modules (which need a context but no weights)."""
import os

_aux_engines = {}


def default_microbatch() -> int:
    """Upper bound on images per internal launch (the library splits a call into equal micro-batches no larger than
    this).  148 = a multiple of 37 = 148 SMs / 4: at 256x256 every layer of the U-Net then has a tile count that is a
    multiple of 148 (e.g. the 32x32 bottleneck: 37 images x 8 tiles x 2 n-tiles = 4 x 148), so the persistent conv
    kernel has no partial last wave; larger launches also amortise the per-layer ramp-up/tail (measured on the
    headline workload: 74 -> 24.0k, 148 -> 24.3k, 222 -> 24.5k images/s).  Workspace: about 53 MB per image at 256x256."""
    return int(os.environ.get("UB_MICROBATCH", "148"))


def new_engine(device, arch, in_channels, n_classes, max_microbatch):
    from .engine import Engine
    return Engine(device, arch, in_channels, n_classes, max_microbatch)


def engine_for(device):
    """Weight-less context for ub_perturb_images / ub_metrics on `device`."""
    import torch
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if idx not in _aux_engines:
        _aux_engines[idx] = new_engine(torch.device("cuda", idx), "unet", 3, 3, 1)
    return _aux_engines[idx]
