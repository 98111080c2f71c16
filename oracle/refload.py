"""Test infrastructure ONLY: import the real reference (MattiDeBeer/image-segmentation) from
/root/reference so the oracle restatements can be pinned against it and golden vectors generated.

/root/reference exists only in the build container, never on the GPU box; nothing in the product
path, `-m gpu` tests, smoke() or bench.py may call this module.  kornia, segmentation_models_pytorch
and matplotlib are absent from the image, and the reference imports them at module top
(models/processing_blocks.py:16-18, models/losses.py:1, models/helperFunctions.py:2,6), so empty
stand-in modules are registered for those names before the import.
"""
import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("UB_REFERENCE_ROOT", "/root/reference")

_STUBS = [
    "kornia", "kornia.augmentation", "kornia.filters",
    "segmentation_models_pytorch", "segmentation_models_pytorch.losses",
    "matplotlib", "matplotlib.pyplot", "matplotlib.patches",
]


def available(root: str = None) -> bool:
    return os.path.isdir(os.path.join(root or REFERENCE_ROOT, "models"))


def _install_stubs():
    for name in _STUBS:
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
            continue
        except Exception:
            pass
        mod = types.ModuleType(name)
        mod.__dict__["__stub__"] = True
        sys.modules[name] = mod
        if "." in name:
            parent, child = name.rsplit(".", 1)
            setattr(sys.modules[parent], child, mod)
    mp = sys.modules["matplotlib.patches"]
    if getattr(mp, "__stub__", False) and not hasattr(mp, "Patch"):
        mp.Patch = type("Patch", (), {})  # models/helperFunctions.py:6 imports the name; plotting is never called
    smp = sys.modules["segmentation_models_pytorch"]
    if getattr(smp, "__stub__", False):
        # models/losses.py:9,22,96,108 instantiate smp.losses.DiceLoss at construction time.
        class _AbsentDiceLoss:  # noqa: D401 - placeholder, never evaluated
            def __init__(self, *a, **k):
                pass

            def __call__(self, *a, **k):
                raise RuntimeError("segmentation_models_pytorch is absent; use oracle.metrics_oracle.dice_smp")
        smp.losses.DiceLoss = _AbsentDiceLoss


def load(root: str = None):
    """Returns a namespace with the reference's own classes (unmodified code).  `root`: /root/reference (build
    container, default) or the byte-for-byte copy under baseline/_ref/ that travels to the GPU box (bench.py's CPU
    legs only; baseline/install_ref.py makes it)."""
    root = root or REFERENCE_ROOT
    if not available(root):
        raise RuntimeError(f"reference tree not found at {root}")
    _install_stubs()
    # The reference uses top-level package names `models` / `customDatasets`; make sure ours
    # (none at repo root) cannot shadow them.
    if root not in sys.path:
        sys.path.insert(0, root)
    ns = types.SimpleNamespace()
    ns.UNet_mod = importlib.import_module("models.UNet")
    ns.blocks = importlib.import_module("models.processing_blocks")
    ns.losses = importlib.import_module("models.losses")
    ns.perturbations = importlib.import_module("customDatasets.perturbations")
    ns.autoencoder = importlib.import_module("models.autoencoder")
    ns.pre_trained = importlib.import_module("models.pre_trained")
    ns.UNet = ns.UNet_mod.UNet
    ns.LargeUNet = ns.UNet_mod.LargeUNet
    return ns


def load_clip_unet(clip_seed: int = 0):
    """The reference's `ClipUnet` class (models/CLIP_models.py:65-134) with the ONE substitution this container forces:
    `CLIPModel.from_pretrained("openai/clip-vit-base-patch32")` (models/processing_blocks.py:209) needs the network, so
    the name `CLIPModel` inside the reference module is bound to a subclass whose `from_pretrained` builds the same
    architecture (default CLIPConfig = ViT-B/32, projection 512) with seeded random weights, and whose
    `get_image_features` returns the tensor the reference's pinned transformers 4.x returned (5.x wraps it in a
    ModelOutput).  Everything else — preprocessing, fusion, U-Net — is the reference's unmodified code."""
    import torch
    import transformers
    ns = load()

    class SeededCLIPModel(transformers.CLIPModel):
        @classmethod
        def from_pretrained(cls, *args, **kwargs):
            torch.manual_seed(clip_seed)
            return cls(transformers.CLIPConfig()).eval()

        def get_image_features(self, *args, **kwargs):
            out = super().get_image_features(*args, **kwargs)
            return out if torch.is_tensor(out) else out.pooler_output

    ns.blocks.CLIPModel = SeededCLIPModel
    ns.CLIP_models = importlib.import_module("models.CLIP_models")
    ns.SeededCLIPModel = SeededCLIPModel
    ns.ClipUnet = ns.CLIP_models.ClipUnet
    # models/prompt_segmentation.py star-imports processing_blocks after the rebinding above, so its
    # ClipFeatureExtractor builds the same seeded stand-in encoder
    ns.prompt_segmentation = importlib.import_module("models.prompt_segmentation")
    ns.ClipUnetPrompt = ns.prompt_segmentation.ClipUnetPrompt
    return ns
