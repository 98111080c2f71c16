"""Drop-in for the parts of the reference's models/processing_blocks.py that the robustness-evaluation
path uses: the four conv blocks (:21-133) and the six float-space perturbation modules (:454-592).

The conv blocks are PARAMETER CONTAINERS with the reference's sub-module names, shapes and default
initialisation, so checkpoints (`state_dict` keys such as `enc1.block.0.conv.0.weight`) load unchanged.
They do not compute on their own: the owning network (models/UNet.py) runs all of them in one pass of
the sm_100a library.  The perturbation modules call the library's perturbation kernels.
"""
from __future__ import annotations

import random

import torch
import torch.nn as nn

from .. import runtime
from ..engine import Perturbation


def _no_standalone(name):
    raise RuntimeError(
        f"{name} is a parameter container in unet-b200; run it through UNet/LargeUNet.forward "
        "(the fused sm_100a path). There is no per-block PyTorch fallback.")


class ConvBlock(nn.Module):
    """(conv3x3 pad1 -> BatchNorm2d -> ReLU) x 2; reference processing_blocks.py:21-52."""

    def __init__(self, in_channels, out_channels, kernel_size=3, padding=1):
        super().__init__()
        if kernel_size != 3 or padding != 1:
            raise ValueError("unet-b200 implements the reference's 3x3 / padding 1 block only")
        self.conv = nn.Sequential(
            nn.Conv2d(in_channels, out_channels, kernel_size, padding=padding),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
            nn.Conv2d(out_channels, out_channels, kernel_size, padding=padding),
            nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
        )

    def forward(self, x):
        _no_standalone("ConvBlock")


class ConvBlockDownsample(nn.Module):
    """ConvBlock followed by MaxPool2d(2,2); reference processing_blocks.py:54-77."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.block = nn.Sequential(ConvBlock(in_channels, out_channels), nn.MaxPool2d(kernel_size=2, stride=2))

    def forward(self, x):
        _no_standalone("ConvBlockDownsample")


class ConvBlockUpsampleSkip(nn.Module):
    """ConvTranspose2d(k2,s2) -> bilinear resize to the skip -> concat -> ConvBlock; reference :79-109."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.up = nn.ConvTranspose2d(in_channels, out_channels, kernel_size=2, stride=2)
        self.conv = ConvBlock(out_channels * 2, out_channels)

    def forward(self, x, skip):
        _no_standalone("ConvBlockUpsampleSkip")


class ConvBlockUpsample(nn.Module):
    """ConvTranspose2d(k2,s2) -> ConvBlock (no skip); reference :111-133.  Container only."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.up = nn.ConvTranspose2d(in_channels, out_channels, kernel_size=2, stride=2)
        self.conv = ConvBlock(out_channels, out_channels)

    def forward(self, x):
        _no_standalone("ConvBlockUpsample")


# --------------------------------------------------------------------------- perturbation modules
class _DevicePerturbation(nn.Module):
    """Common part: `forward(img[B,3,H,W] fp32 on cuda) -> img`, computed by ub_perturb_images.

    `spec(img)` returns the engine.Perturbation for this call; random families draw their numbers
    exactly where the reference draws them (so seeding `random` / `torch` reproduces its output)."""

    def spec(self, img) -> Perturbation:
        raise NotImplementedError

    def forward(self, img):
        if not img.is_cuda:
            raise RuntimeError(f"{type(self).__name__}: unet-b200 perturbations run on CUDA only")
        eng = runtime.engine_for(img.device)
        return eng.perturb(img.float().contiguous(), self.spec(img))


class GaussianPixelNoise(_DevicePerturbation):
    """clamp(img + randn*std/255, 0, 1); reference :454-474.  The normal draw stays `torch.randn_like`
    (same generator, same order as the reference); the arithmetic runs in the CUDA kernel."""

    def __init__(self, std):
        super().__init__()
        self.std = std

    def spec(self, img):
        return Perturbation.f_noise(self.std, aux=torch.randn_like(img))


class RepeatedBlur(_DevicePerturbation):
    """`times` passes of kornia.filters.box_blur(img,(3,3)) (reflect border); reference :477-496."""

    def __init__(self, times):
        super().__init__()
        self.times = times

    def spec(self, img):
        return Perturbation.f_blur(self.times)


class ContrastChange(_DevicePerturbation):
    """clamp(img*factor, 0, 1); reference :499-518."""

    def __init__(self, factor):
        super().__init__()
        self.factor = factor

    def spec(self, img):
        return Perturbation.f_contrast(self.factor)


class BrightnessChange(_DevicePerturbation):
    """clamp(img + offset/255, 0, 1); reference :521-539."""

    def __init__(self, offset):
        super().__init__()
        self.offset = offset / 255.0

    def spec(self, img):
        p = Perturbation.f_brightness(0.0)
        p.struct.fparam = float(self.offset)
        return p


class Occlusion(_DevicePerturbation):
    """Zero a size x size square per image, corner from Python `random` (x drawn before y);
    reference :542-563.  Like the reference it writes INTO its input tensor and returns it."""

    def __init__(self, size):
        super().__init__()
        self.size = size

    def corners(self, b, h, w):
        out = []
        for _ in range(b):
            x = random.randint(0, w - self.size) if w > self.size else 0
            y = random.randint(0, h - self.size) if h > self.size else 0
            out.append((y, x))
        return torch.tensor(out, dtype=torch.int32).reshape(b, 2)

    def spec(self, img):
        b, _, h, w = img.shape
        return Perturbation.f_occlusion(self.size, aux=self.corners(b, h, w))

    def forward(self, img):
        out = super().forward(img)
        if img.dtype == torch.float32 and img.is_contiguous():
            img.copy_(out)
            return img
        return out


class SaltAndPepper(_DevicePerturbation):
    """One uniform draw per pixel: < amount/2 -> 1.0, > 1-amount/2 -> 0.0; reference :565-592."""

    def __init__(self, amount):
        super().__init__()
        self.amount = amount

    def spec(self, img):
        b, _, h, w = img.shape
        return Perturbation.f_salt_pepper(self.amount, aux=torch.rand((b, 1, h, w), device=img.device))


# --------------------------------------------------------------------------- ClipUnet blocks (reference :135-233, :286-322)
class CustomClipPreprocessor(nn.Module):
    """Resize to 224x224 and normalise with the CLIP statistics; reference :135-168.  The reference loops over the
    images of the batch; torchvision's Resize / Normalize act on the trailing [C,H,W] dimensions, so one batched call
    gives the same values."""

    def __init__(self, mean, std, target_size=(224, 224)):
        super().__init__()
        from torchvision import transforms
        self.resize = transforms.Resize(target_size)
        self.normalize = transforms.Normalize(mean=mean, std=std)

    def forward(self, images):
        return self.normalize(self.resize(images))


class ClipFeatureExtractor(nn.Module):
    """CLIP image embedding [B,E]; reference :171-233 (`CustomClipPreprocessor` + `CLIPModel.get_image_features`).

    backend "b200" (default): Resize + Normalize + the ViT image tower + the visual projection all run in the library
    (`ub_clip_image_features`: bf16 tensor-core GEMMs with fp32 accumulation, fp32 residual stream / LayerNorm / softmax);
    the `clip_model` sub-module only holds the parameters (same state_dict keys as the reference, so checkpoints load).
    backend "transformers": HuggingFace's CLIPModel on PyTorch kernels exactly as the reference runs it (fp32, or bf16
    under CUDA autocast when `autocast_dtype` is given) — an explicit choice, never a silent fallback.
    `clip_model` lets a caller supply an already constructed CLIPModel (e.g. random-init where the
    `openai/clip-vit-base-patch32` checkpoint cannot be downloaded); the default loads the checkpoint like the
    reference (:209)."""

    def __init__(self, train=False, clip_model=None, autocast_dtype=None, backend=None):
        super().__init__()
        if backend is None:
            backend = "transformers" if autocast_dtype is not None else "b200"
        if backend not in ("b200", "transformers"):
            raise ValueError(f"unknown CLIP backend {backend!r}")
        self.backend = backend
        self.autocast_dtype = autocast_dtype
        self.device = torch.device('cuda' if torch.cuda.is_available() else 'cpu')
        mean = [0.48145466, 0.4578275, 0.40821073]
        std = [0.26862954, 0.26130258, 0.27577711]
        self.custom_preprocessor = CustomClipPreprocessor(mean=mean, std=std)
        if clip_model is None:
            from transformers import CLIPModel
            clip_model = CLIPModel.from_pretrained("openai/clip-vit-base-patch32")
        self.clip_model = clip_model.to(self.device)
        self.train_clip = train
        self.set_train(train)
        self._vit_engines = {}  # (device, max_microbatch) -> (weights key, Engine)
        self._vit_params = None

    def set_train(self, value: bool):
        assert isinstance(value, bool), "Value must be a boolean"
        if value:
            raise RuntimeError("unet-b200 is inference-only: the CLIP encoder stays frozen")
        for param in self.clip_model.parameters():
            param.requires_grad = value
        self.train_clip = value

    def _check_supported(self):
        cfg = getattr(getattr(self.clip_model, "config", None), "vision_config", None)
        if cfg is None:
            raise RuntimeError("the b200 CLIP backend needs a HuggingFace CLIPModel (config.vision_config); "
                               "pass backend='transformers' for anything else")
        bad = []
        if cfg.hidden_act != "quick_gelu":
            bad.append(f"hidden_act {cfg.hidden_act!r} (quick_gelu)")
        if abs(cfg.layer_norm_eps - 1e-5) > 1e-12:
            bad.append(f"layer_norm_eps {cfg.layer_norm_eps} (1e-5)")
        if cfg.num_attention_heads * 64 != cfg.hidden_size:
            bad.append(f"head width {cfg.hidden_size // cfg.num_attention_heads} (64)")
        if cfg.num_channels != 3:
            bad.append(f"{cfg.num_channels} channels (3)")
        if bad:
            raise RuntimeError("CLIP vision tower not supported by the b200 backend: " + ", ".join(bad) +
                               "; pass backend='transformers'")

    def vit_engine(self, device, max_microbatch=None):
        """The library context holding the packed vision tower (rebuilt when the weights change)."""
        from .. import runtime
        dev = torch.device(device)
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        mb = max_microbatch or runtime.default_microbatch()
        if self._vit_params is None:
            self._check_supported()
            self._vit_params = [p for n, p in self.clip_model.named_parameters()
                                if n.startswith("vision_model.") or n == "visual_projection.weight"]
        wkey = tuple((p.data_ptr(), p._version) for p in self._vit_params)
        slot = self._vit_engines.get((dev, mb))
        if slot is None or slot[0] != wkey:
            if slot is not None:
                slot[1].close("the CLIP weights changed and the engine was rebuilt")
            eng = runtime.new_engine(dev, "clip_vit", 3, 1, mb)
            eng.load_state_dict({k: v for k, v in self.clip_model.state_dict().items()
                                 if k.startswith("vision_model.") or k == "visual_projection.weight"})
            self._vit_engines[(dev, mb)] = (wkey, eng)
            slot = self._vit_engines[(dev, mb)]
        return slot[1]

    def forward(self, X):
        if self.backend == "b200":
            if not X.is_cuda:
                raise RuntimeError("unet-b200 has no CPU path: inputs must be CUDA tensors")
            return self.vit_engine(X.device).clip_image_features(X)
        inputs = self.custom_preprocessor(X)
        with torch.no_grad(), torch.autocast("cuda", dtype=self.autocast_dtype or torch.bfloat16,
                                             enabled=self.autocast_dtype is not None and inputs.is_cuda):
            feats = self.clip_model.get_image_features(pixel_values=inputs)
        if not torch.is_tensor(feats):  # transformers >= 5 returns a ModelOutput; the projected embedding is its pooler_output
            feats = feats.pooler_output
        return feats.float()


class CrossAttentionFusion(nn.Module):
    """nn.MultiheadAttention(query = bottleneck positions, key = value = the CLIP embedding); reference :286-322.
    Parameter container: with one key per image the attention weights are identically 1, so the library evaluates
    out_proj(v_proj(clip)) once per image and broadcasts it (csrc/ub_aux.cuh: clip_fusion_broadcast_kernel)."""

    def __init__(self, resnet_channels, num_heads=4):
        super().__init__()
        self.cross_attn = nn.MultiheadAttention(embed_dim=resnet_channels, num_heads=num_heads)

    def forward(self, resnet_feats, clip_feats):
        _no_standalone("CrossAttentionFusion")
