"""Host-side handle on one ub_ctx (one per device per process).  PyTorch supplies device memory and the
stream; every computation is a call into libunetb200.so through the C ABI (include/unet_b200.h)."""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import torch

from . import _lib as L


class Perturbation:
    """Host description of one (family, level): the ub_perturb struct plus an optional aux tensor
    (reference-generated noise / corners / code map in parity mode; None = device RNG)."""

    def __init__(self, kind: int = L.UB_P_NONE, iparam: int = 0, fparam: float = 0.0, fparam2: float = 0.0,
                 seed: int = 0, aux: Optional[torch.Tensor] = None, lut=None, name: str = "none"):
        self.struct = L.ub_perturb()
        self.struct.kind = kind
        self.struct.iparam = int(iparam)
        self.struct.fparam = float(fparam)
        self.struct.fparam2 = float(fparam2)
        self.struct.seed = int(seed) & (2 ** 64 - 1)
        if lut is not None:
            lib = L.load()
            rc = lib.ub_make_lut(ctypes.byref(self.struct), int(lut[0]), float(lut[1]))
            if rc != 0:
                raise ValueError(f"ub_make_lut failed ({rc})")
        self.aux = aux
        self.name = name

    # ---- uint8-space Dataset-wrapper families (customDatasets/perturbations.py)
    @staticmethod
    def u8_noise(std, aux=None, seed=0):
        return Perturbation(L.UB_P_U8_NOISE, fparam=float(std), aux=aux, seed=seed, name="gaussian_noise")

    @staticmethod
    def u8_blur(passes):
        return Perturbation(L.UB_P_U8_BLUR, iparam=int(passes), name="gaussian_blur")

    @staticmethod
    def u8_contrast(scale):
        return Perturbation(lut=(L.UB_LUT_CONTRAST, scale), name="contrast")

    @staticmethod
    def u8_brightness(offset, increase=True):
        return Perturbation(lut=(L.UB_LUT_BRIGHTNESS_INCREASE if increase else L.UB_LUT_BRIGHTNESS_DECREASE, offset),
                            name="brightness")

    @staticmethod
    def u8_occlusion(size, aux=None, seed=0):
        return Perturbation(L.UB_P_U8_OCCLUSION, iparam=int(size), aux=aux, seed=seed, name="occlusion")

    @staticmethod
    def u8_salt_pepper(amount, aux=None, seed=0, hw=256 * 256):
        n = int(round(amount * hw))
        p_hit = 1.0 - (1.0 - 1.0 / hw) ** n  # P(pixel drawn at least once), perturbations.py:366-377
        return Perturbation(L.UB_P_U8_SALT_PEPPER, fparam=float(amount), fparam2=p_hit, aux=aux, seed=seed,
                            name="salt_pepper_noise")

    # ---- float-space nn.Module families (models/processing_blocks.py:454-592)
    @staticmethod
    def f_noise(std, aux=None, seed=0):
        return Perturbation(L.UB_P_F_NOISE, fparam=float(std), fparam2=std / 255.0, aux=aux, seed=seed,
                            name="GaussianPixelNoise")

    @staticmethod
    def f_blur(times):
        return Perturbation(L.UB_P_F_BLUR, iparam=int(times), name="RepeatedBlur")

    @staticmethod
    def f_contrast(factor):
        return Perturbation(L.UB_P_F_CONTRAST, fparam=float(factor), name="ContrastChange")

    @staticmethod
    def f_brightness(offset):
        return Perturbation(L.UB_P_F_BRIGHTNESS, fparam=offset / 255.0, name="BrightnessChange")

    @staticmethod
    def f_occlusion(size, aux=None, seed=0):
        return Perturbation(L.UB_P_F_OCCLUSION, iparam=int(size), aux=aux, seed=seed, name="Occlusion")

    @staticmethod
    def f_salt_pepper(amount, aux=None, seed=0):
        return Perturbation(L.UB_P_F_SALT_PEPPER, fparam=amount / 2, fparam2=1 - amount / 2, aux=aux, seed=seed,
                            name="SaltAndPepper")


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Engine:
    """Owns one ub_ctx.  arch: 'unet' | 'large' | 'clip' (ClipUnet: every evaluation call also takes the CLIP
    image embeddings of the same images, `clip_feats` fp32 [n, 512]) | 'clip_prompt' (ClipUnetPrompt: forward_prompt
    takes the embeddings and the 1-channel prompt heat-maps) | 'autoencoder' (forward returns the sigmoid reconstruction;
    n_classes = its output channels)."""

    def __init__(self, device, arch: str = "unet", in_channels: int = 3, n_classes: int = 3,
                 max_microbatch: int = 148):
        self.lib = L.load()
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("unet-b200 runs on CUDA sm_100a devices only (no CPU fallback)")
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", idx)
        self.n_classes = n_classes
        self.max_microbatch = max_microbatch
        h = ctypes.c_void_p()
        archs = {"unet": L.UB_ARCH_UNET, "large": L.UB_ARCH_LARGE_UNET, "clip": L.UB_ARCH_CLIP_UNET,
                 "clip_prompt": L.UB_ARCH_CLIP_UNET_PROMPT, "autoencoder": L.UB_ARCH_AUTOENCODER,
                 "encoder": L.UB_ARCH_ENCODER, "skip_decoder": L.UB_ARCH_SKIP_DECODER, "clip_vit": L.UB_ARCH_CLIP_VIT}
        if arch not in archs:
            raise ValueError(f"unknown arch {arch!r}")
        self.arch = arch
        rc = self.lib.ub_create(ctypes.byref(h), idx, archs[arch], in_channels, n_classes, max_microbatch)
        if rc != 0:
            raise RuntimeError(f"ub_create failed ({rc}): {self.lib.ub_last_error(None).decode()}")
        self.h = h
        self.loaded = False

    def close(self, why: str = "this engine has been closed"):
        if getattr(self, "h", None):
            self.lib.ub_destroy(self.h)
            self.h = None
            self._closed_why = why

    def _handle(self):
        if not getattr(self, "h", None):
            raise RuntimeError(getattr(self, "_closed_why", "this engine has been closed"))
        return self.h

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"{what} failed ({rc}): {self.lib.ub_last_error(self.h).decode()}")

    def load_state_dict(self, state_dict):
        items = [(k, v.detach().to("cpu", torch.float32).contiguous()) for k, v in state_dict.items()
                 if torch.is_tensor(v) and v.is_floating_point()]
        n = len(items)
        names = (ctypes.c_char_p * n)(*[k.encode() for k, _ in items])
        ptrs = (ctypes.c_void_p * n)(*[v.data_ptr() for _, v in items])
        shape_arrays = [(ctypes.c_int64 * max(v.dim(), 1))(*(list(v.shape) or [1])) for _, v in items]
        shapes = (ctypes.POINTER(ctypes.c_int64) * n)(*[ctypes.cast(a, ctypes.POINTER(ctypes.c_int64))
                                                        for a in shape_arrays])
        ndims = (ctypes.c_int * n)(*[v.dim() for _, v in items])
        self._check(self.lib.ub_load_weights(self._handle(), n, names, ptrs, shapes, ndims), "ub_load_weights")
        self.loaded = True

    def forward_prompt(self, x: torch.Tensor, prompt_heatmap: torch.Tensor, clip_feats: torch.Tensor) -> torch.Tensor:
        """ClipUnetPrompt.forward given the CLIP embedding (models/prompt_segmentation.py:63-95): -> logits [n,C,H,W]."""
        if self.arch != "clip_prompt":
            raise RuntimeError("forward_prompt needs an engine created with arch='clip_prompt'")
        x = x.to(self.device, torch.float32).contiguous()
        n, c, h, w = x.shape
        hm = prompt_heatmap.to(self.device, torch.float32).contiguous()
        if tuple(hm.shape) != (n, 1, h, w):
            raise ValueError(f"prompt_heatmap must be [{n}, 1, {h}, {w}], got {tuple(hm.shape)}")
        cf = clip_feats.to(self.device, torch.float32).contiguous()
        if tuple(cf.shape) != (n, 512):
            raise ValueError(f"clip_feats must be [{n}, 512], got {tuple(cf.shape)}")
        out = torch.empty((n, self.n_classes, h, w), device=self.device, dtype=torch.float32)
        self._check(self.lib.ub_forward_prompt(self._handle(), _ptr(x), _ptr(hm), _ptr(cf), _ptr(out), n, h, w,
                                               _stream(self.device)), "ub_forward_prompt")
        return out

    FEATURE_NAMES = ("x0", "enc1", "enc2", "enc3", "bottleneck")  # models/autoencoder.py:25

    def encoder_forward(self, x: torch.Tensor) -> dict:
        """Encoder.forward (models/autoencoder.py:17-25): x fp32 [n,3,H,W] -> {"x0","enc1","enc2","enc3","bottleneck"}."""
        if self.arch != "encoder":
            raise RuntimeError("encoder_forward needs an engine created with arch='encoder'")
        x = x.to(self.device, torch.float32).contiguous()
        n, _, h, w = x.shape
        shapes = [(32, 0), (64, 1), (128, 2), (256, 3), (512, 3)]
        feats = [torch.empty((n, c, h >> l, w >> l), device=self.device, dtype=torch.float32) for c, l in shapes]
        ptrs = (ctypes.c_void_p * len(feats))(*[t.data_ptr() for t in feats])
        self._check(self.lib.ub_encoder_forward(self._handle(), _ptr(x), n, h, w, ptrs, len(feats), _stream(self.device)),
                    "ub_encoder_forward")
        return dict(zip(self.FEATURE_NAMES, feats))

    def skip_decoder_forward(self, features: dict) -> torch.Tensor:
        """SegmentationDecoderSkip.forward (models/pre_trained.py:17-24): the encoder's dict -> logits fp32 [n,C,H,W]."""
        if self.arch != "skip_decoder":
            raise RuntimeError("skip_decoder_forward needs an engine created with arch='skip_decoder'")
        feats = [features[k].to(self.device, torch.float32).contiguous() for k in self.FEATURE_NAMES]
        n, _, h, w = feats[0].shape
        for t, (c, l) in zip(feats, [(32, 0), (64, 1), (128, 2), (256, 3), (512, 3)]):
            if tuple(t.shape) != (n, c, h >> l, w >> l):
                raise ValueError(f"feature map of shape {tuple(t.shape)}; expected {(n, c, h >> l, w >> l)}")
        out = torch.empty((n, self.n_classes, h, w), device=self.device, dtype=torch.float32)
        ptrs = (ctypes.c_void_p * len(feats))(*[t.data_ptr() for t in feats])
        self._check(self.lib.ub_skip_decoder_forward(self._handle(), ptrs, len(feats), _ptr(out), n, h, w,
                                                     _stream(self.device)), "ub_skip_decoder_forward")
        return out

    def forward(self, x: torch.Tensor, clip_feats: Optional[torch.Tensor] = None) -> torch.Tensor:
        x = x.to(self.device, torch.float32).contiguous()
        n, c, h, w = x.shape
        if self.arch == "clip":
            return self.eval_batches(x, None, None, 1, want_logits=True, clip_feats=clip_feats)["logits"]
        out = torch.empty((n, self.n_classes, h, w), device=self.device, dtype=torch.float32)
        self._check(self.lib.ub_forward(self._handle(), _ptr(x), _ptr(out), n, h, w, _stream(self.device)), "ub_forward")
        return out

    def eval_batches(self, images: torch.Tensor, targets: Optional[torch.Tensor], perturbation: Optional[Perturbation],
                     ref_batch: int, want_dice: bool = False, want_pred: bool = False, want_logits: bool = False,
                     want_perturbed: bool = False, out_cm: Optional[torch.Tensor] = None,
                     out_dice: Optional[torch.Tensor] = None, clip_feats: Optional[torch.Tensor] = None):
        """images: uint8 [n,H,W,3] or fp32 [n,3,H,W]; targets: uint8 or int64 [n,H,W] (device tensors).
        Returns dict(cm=int64 [nb,C,C], dice=f64 [nb,C,3], pred=u8 [n,H,W], logits, perturbed)."""
        assert images.is_cuda and images.is_contiguous()
        if images.dtype == torch.uint8:
            fmt = L.UB_IMG_U8_NHWC
            n, h, w, _ = images.shape
        elif images.dtype == torch.float32:
            fmt = L.UB_IMG_F32_NCHW
            n, _, h, w = images.shape
        else:
            raise TypeError("images must be uint8 NHWC or float32 NCHW")
        C = self.n_classes
        nb = math.ceil(n / ref_batch)
        tfmt = L.UB_TGT_I64
        cm = dice = None
        if targets is not None:
            assert targets.is_cuda and targets.is_contiguous() and tuple(targets.shape) == (n, h, w)
            if targets.dtype == torch.uint8:
                tfmt = L.UB_TGT_U8
            elif targets.dtype != torch.int64:
                raise TypeError("targets must be uint8 or int64")
            cm = out_cm if out_cm is not None else torch.empty((nb, C, C), device=self.device, dtype=torch.int64)
            if want_dice:
                dice = out_dice if out_dice is not None else torch.empty((nb, C, 3), device=self.device,
                                                                         dtype=torch.float64)
        pred = torch.empty((n, h, w), device=self.device, dtype=torch.uint8) if want_pred else None
        logits = torch.empty((n, C, h, w), device=self.device, dtype=torch.float32) if want_logits else None
        pert = torch.empty((n, 3, h, w), device=self.device, dtype=torch.float32) if want_perturbed else None
        p = perturbation
        aux = None
        if p is not None and p.aux is not None:
            aux = p.aux.to(self.device).contiguous()
        if self.arch == "clip":
            if clip_feats is None:
                raise ValueError("a ClipUnet engine needs clip_feats (fp32 [n, 512], the CLIP image embeddings)")
            clip_feats = clip_feats.to(self.device, torch.float32).contiguous()
            if tuple(clip_feats.shape) != (n, 512):
                raise ValueError(f"clip_feats must be [n, 512], got {tuple(clip_feats.shape)}")
            self._check(self.lib.ub_eval_batches_clip(
                self._handle(), ctypes.byref(p.struct) if p is not None else None, _ptr(images), fmt, _ptr(targets), tfmt,
                _ptr(aux), _ptr(clip_feats), n, h, w, int(ref_batch), _ptr(cm), _ptr(dice), _ptr(pred), _ptr(logits),
                _ptr(pert), _stream(self.device)), "ub_eval_batches_clip")
            return dict(cm=cm, dice=dice, pred=pred, logits=logits, perturbed=pert)
        self._check(self.lib.ub_eval_batches(self._handle(), ctypes.byref(p.struct) if p is not None else None, _ptr(images),
                                             fmt, _ptr(targets), tfmt, _ptr(aux), n, h, w, int(ref_batch), _ptr(cm),
                                             _ptr(dice), _ptr(pred), _ptr(logits), _ptr(pert),
                                             _stream(self.device)), "ub_eval_batches")
        return dict(cm=cm, dice=dice, pred=pred, logits=logits, perturbed=pert)

    def perturb(self, images: torch.Tensor, perturbation: Perturbation) -> torch.Tensor:
        assert images.is_cuda and images.is_contiguous()
        if images.dtype == torch.uint8:
            fmt = L.UB_IMG_U8_NHWC
            n, h, w, _ = images.shape
        else:
            fmt = L.UB_IMG_F32_NCHW
            n, _, h, w = images.shape
        out = torch.empty((n, 3, h, w), device=self.device, dtype=torch.float32)
        aux = perturbation.aux.to(self.device).contiguous() if perturbation.aux is not None else None
        self._check(self.lib.ub_perturb_images(self._handle(), ctypes.byref(perturbation.struct), _ptr(images), fmt, _ptr(aux),
                                               n, h, w, _ptr(out), _stream(self.device)), "ub_perturb_images")
        return out

    def clip_dims(self):
        """(input size S, embedding width E, images per internal pass) of a loaded 'clip_vit' engine."""
        if getattr(self, "_clip_dims", None) is None:
            s, e, c = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
            self._check(self.lib.ub_clip_dims(self._handle(), ctypes.byref(s), ctypes.byref(e), ctypes.byref(c)), "ub_clip_dims")
            self._clip_dims = (s.value, e.value, c.value)
        return self._clip_dims

    def clip_image_features(self, x: torch.Tensor) -> torch.Tensor:
        """ClipFeatureExtractor.forward (models/processing_blocks.py:219-233): fp32 [n,3,H,W] -> CLIP image embeddings
        fp32 [n,E]; Resize + Normalize + ViT + projection, all in the library ('clip_vit' engines)."""
        if self.arch != "clip_vit":
            raise RuntimeError("clip_image_features needs an engine created with arch='clip_vit'")
        x = x.to(self.device, torch.float32).contiguous()
        n, c, h, w = x.shape
        if c != 3:
            raise ValueError("CLIP takes 3-channel images")
        out = torch.empty((n, self.clip_dims()[1]), device=self.device, dtype=torch.float32)
        self._check(self.lib.ub_clip_image_features(self._handle(), _ptr(x), n, h, w, _ptr(out), _stream(self.device)),
                    "ub_clip_image_features")
        return out

    def clip_preprocess(self, x: torch.Tensor) -> torch.Tensor:
        """CustomClipPreprocessor.forward alone (test hook): fp32 [n,3,H,W] -> fp32 [n,3,S,S]."""
        x = x.to(self.device, torch.float32).contiguous()
        n, _, h, w = x.shape
        s = self.clip_dims()[0]
        out = torch.empty((n, 3, s, s), device=self.device, dtype=torch.float32)
        self._check(self.lib.ub_debug_clip_preprocess(self._handle(), _ptr(x), n, h, w, _ptr(out), _stream(self.device)),
                    "ub_debug_clip_preprocess")
        return out

    def blur_static(self, images: torch.Tensor, perturbation: Perturbation, out: torch.Tensor) -> torch.Tensor:
        """`ub_blur_images_static`: the blur families with the library's static scratch planes — no device allocation in
        the call, so the CALLER provides `out` (fp32 [n,3,H,W]).  No counterpart in the reference (include/unet_b200.h)."""
        assert images.is_cuda and images.is_contiguous() and out.is_cuda and out.is_contiguous() and out.dtype == torch.float32
        if images.dtype == torch.uint8:
            fmt = L.UB_IMG_U8_NHWC
            n, h, w, _ = images.shape
        else:
            fmt = L.UB_IMG_F32_NCHW
            n, _, h, w = images.shape
        assert tuple(out.shape) == (n, 3, h, w)
        self._check(self.lib.ub_blur_images_static(self._handle(), ctypes.byref(perturbation.struct), _ptr(images), fmt,
                                                   n, h, w, _ptr(out), _stream(self.device)), "ub_blur_images_static")
        return out

    def metrics(self, logits: torch.Tensor, targets: torch.Tensor, ref_batch: int, want_dice: bool = False):
        logits = logits.to(self.device, torch.float32).contiguous()
        targets = targets.to(self.device).contiguous()
        n, C, h, w = logits.shape
        tfmt = L.UB_TGT_U8 if targets.dtype == torch.uint8 else L.UB_TGT_I64
        if tfmt == L.UB_TGT_I64 and targets.dtype != torch.int64:
            targets = targets.long()
        nb = math.ceil(n / ref_batch)
        cm = torch.empty((nb, C, C), device=self.device, dtype=torch.int64)
        dice = torch.empty((nb, C, 3), device=self.device, dtype=torch.float64) if want_dice else None
        self._check(self.lib.ub_metrics(self._handle(), _ptr(logits), _ptr(targets), tfmt, n, C, h, w, int(ref_batch), _ptr(cm),
                                        _ptr(dice), _stream(self.device)), "ub_metrics")
        return cm, dice

    def decode_masks(self, raw: torch.Tensor) -> torch.Tensor:
        """Raw trimap bytes uint8 [n, ...] (38 cat / 75 dog / 255 border) -> class ids uint8, same shape; the
        per-image rule of customDatasets/datasets.py:118-131."""
        if raw.dtype != torch.uint8 or raw.dim() < 2:
            raise ValueError("decode_masks expects a uint8 tensor [n, ...]")
        raw = raw.to(self.device).contiguous()
        out = torch.empty_like(raw)
        n = raw.shape[0]
        px = raw[0].numel() if n else 1
        self._check(self.lib.ub_decode_masks(self._handle(), _ptr(raw), _ptr(out), n, px, _stream(self.device)),
                    "ub_decode_masks")
        return out

    def pack_dataset(self, images: Optional[torch.Tensor], masks: Optional[torch.Tensor]):
        """(fp32 [n,3,H,W] in [0,1], int64 [n,H,W]) -> (uint8 [n,H,W,3], uint8 [n,H,W]) on the device; either may
        be None.  Inverse of the reference's `/255` (datasets.py:115) for the k/255 values it yields."""
        src = images if images is not None else masks
        n, h, w = src.shape[0], src.shape[-2], src.shape[-1]
        img_u8 = mask_u8 = None
        if images is not None:
            images = images.to(self.device, torch.float32).contiguous()
            img_u8 = torch.empty((n, h, w, 3), device=self.device, dtype=torch.uint8)
        if masks is not None:
            masks = masks.to(self.device, torch.int64).contiguous()
            mask_u8 = torch.empty((n, h, w), device=self.device, dtype=torch.uint8)
        self._check(self.lib.ub_pack_dataset(self._handle(), _ptr(images), _ptr(masks), n, h, w, _ptr(img_u8), _ptr(mask_u8),
                                             _stream(self.device)), "ub_pack_dataset")
        return img_u8, mask_u8

    @staticmethod
    def _to_blocked(x_nhwc: torch.Tensor) -> torch.Tensor:
        """NHWC -> the library's activation layout [n][C/8][h][w][8] (channel groups of 8 = 16 bytes)."""
        n, h, w, c = x_nhwc.shape
        return x_nhwc.view(n, h, w, c // 8, 8).permute(0, 3, 1, 2, 4).contiguous()

    @staticmethod
    def _from_blocked(x: torch.Tensor) -> torch.Tensor:
        n, g, h, w, _ = x.shape
        return x.permute(0, 2, 3, 1, 4).reshape(n, h, w, g * 8).contiguous()

    def debug_conv3x3(self, key: str, x_nhwc_bf16: torch.Tensor, pool: bool = False) -> torch.Tensor:
        """One conv3x3+BN+ReLU layer on NHWC bf16 data (converted to / from the blocked layout here)."""
        n, h, w, _ = x_nhwc_bf16.shape
        cout = self._layer_cout(key)
        oh, ow = (h // 2, w // 2) if pool else (h, w)
        out = torch.zeros((n, cout // 8, oh, ow, 8), device=self.device, dtype=torch.bfloat16)
        xb = self._to_blocked(x_nhwc_bf16)
        self._check(self.lib.ub_debug_conv3x3(self._handle(), key.encode(), _ptr(xb), _ptr(out), n, h, w,
                                              int(pool), _stream(self.device)), "ub_debug_conv3x3")
        return self._from_blocked(out)

    def debug_convt2x2(self, key: str, x_nhwc_bf16: torch.Tensor) -> torch.Tensor:
        n, h, w, _ = x_nhwc_bf16.shape
        cout = self._layer_cout(key)
        out = torch.zeros((n, cout // 8, 2 * h, 2 * w, 8), device=self.device, dtype=torch.bfloat16)
        xb = self._to_blocked(x_nhwc_bf16)
        self._check(self.lib.ub_debug_convt2x2(self._handle(), key.encode(), _ptr(xb), _ptr(out), n, h, w,
                                               _stream(self.device)), "ub_debug_convt2x2")
        return self._from_blocked(out)

    def debug_resample(self, x_nhwc_bf16: torch.Tensor, ho: int, wo: int) -> torch.Tensor:
        """The decoder's bilinear align_corners=True resize alone (processing_blocks.py:107); NHWC bf16 in and out."""
        n, h, w, c = x_nhwc_bf16.shape
        out = torch.zeros((n, c // 8, ho, wo, 8), device=self.device, dtype=torch.bfloat16)
        xb = self._to_blocked(x_nhwc_bf16.to(self.device))
        self._check(self.lib.ub_debug_resample(self._handle(), _ptr(xb), _ptr(out), n, c, h, w, ho, wo, _stream(self.device)),
                    "ub_debug_resample")
        return self._from_blocked(out)

    def _layer_cout(self, key):
        return self._couts[key]

    def set_layer_couts(self, couts):
        self._couts = dict(couts)

    def profile_begin(self):
        self._check(self.lib.ub_profile_begin(self._handle()), "ub_profile_begin")

    def profile_end(self):
        """-> (gemm_ms_total, gemm_launches, [(key, cin, n_total, ksize, launches, ms), ...])"""
        ms, cnt = ctypes.c_double(), ctypes.c_int64()
        buf = ctypes.create_string_buffer(16384)
        self._check(self.lib.ub_profile_end(self._handle(), ctypes.byref(ms), ctypes.byref(cnt), buf, len(buf)), "ub_profile_end")
        rows = []
        for line in buf.value.decode().splitlines():
            k, cin, nt, ks, n, t = line.split()
            rows.append((k, int(cin), int(nt), int(ks), int(n), float(t)))
        return ms.value, cnt.value, rows

    @property
    def launch_count(self) -> int:
        return int(self.lib.ub_launch_count(self._handle()))
