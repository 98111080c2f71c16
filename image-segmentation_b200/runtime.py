"""Process-wide bits: default micro-batch, and one shared Engine per device for the stand-alone
perturbation / metric modules (which need a context but no weights)."""
import os

_aux_engines = {}


def default_microbatch() -> int:
    return int(os.environ.get("UB_MICROBATCH", "32"))


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
