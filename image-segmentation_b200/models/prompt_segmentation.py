"""Drop-in for the reference's models/prompt_segmentation.py: `PromptEncoder` (:16-29) and `ClipUnetPrompt` (:31-95).

Same constructors, same sub-module tree and state_dict keys.  `ClipUnetPrompt.forward(X, prompt_heatmap)` is
    clip_features = CLIP ViT-B/32 image embedding of X             (HuggingFace transformers, as in the reference)
    logits        = unet-b200 library (ub_forward_prompt): stem + enc1-3 on X, PromptEncoder on the heat-map,
                    cross-attention fusion of clip_features, prompt_fusion (1x1, 1024 -> 512) over their channel
                    concat, dec1-4 with the encoder skips, 1x1 head
As in `ClipUnet`, `self.bottleneck(enc3)` (:73) is dead compute in the reference — `CrossAttentionFusion` attends to a
single CLIP token, so its output does not depend on the queries — and is not evaluated; its parameters are kept so
checkpoints load.  `PromptEncoder` on its own is a parameter container (the library runs it as part of the model).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .processing_blocks import (ClipFeatureExtractor, ConvBlock, ConvBlockDownsample, ConvBlockUpsampleSkip,
                                CrossAttentionFusion)
from .CLIP_models import ClipUnet


class PromptEncoder(nn.Module):
    """1 -> 32 -> 64 -> 128 (three pooled stages) -> out_channels at 1/8 resolution; reference :16-29."""

    def __init__(self, out_channels=512):
        super().__init__()
        self.enc1 = ConvBlockDownsample(1, 32)
        self.enc2 = ConvBlockDownsample(32, 64)
        self.enc3 = ConvBlockDownsample(64, 128)
        self.conv = ConvBlock(128, out_channels)

    def forward(self, x):
        raise RuntimeError("unet-b200 runs PromptEncoder inside ClipUnetPrompt (ub_forward_prompt); it has no "
                           "stand-alone path")


class ClipUnetPrompt(ClipUnet):
    _arch = "clip_prompt"

    def __init__(self, out_channels=1, in_channels=3, activation=nn.Identity(), clip_model=None, clip_autocast_dtype=None, clip_backend=None):
        nn.Module.__init__(self)
        # registration order follows the reference's __init__ (:33-61), so a seeded construction draws the same weights
        self.clip_feature_extractor = ClipFeatureExtractor(train=False, clip_model=clip_model,
                                                           autocast_dtype=clip_autocast_dtype, backend=clip_backend)
        self.cross_attention_fusion = CrossAttentionFusion(512, num_heads=1)
        self.input = nn.Conv2d(in_channels, 32, kernel_size=1, padding=0)
        self.enc1 = ConvBlockDownsample(32, 64)
        self.enc2 = ConvBlockDownsample(64, 128)
        self.enc3 = ConvBlockDownsample(128, 256)
        self.bottleneck = ConvBlock(256, 512)
        self.prompt_encoder = PromptEncoder(out_channels=512)
        self.prompt_fusion = nn.Conv2d(1024, 512, kernel_size=1, padding=0)
        self.dec1 = ConvBlockUpsampleSkip(512, 256)
        self.dec2 = ConvBlockUpsampleSkip(256, 128)
        self.dec3 = ConvBlockUpsampleSkip(128, 64)
        self.dec4 = ConvBlockUpsampleSkip(64, 32)
        self.out = nn.Conv2d(32, out_channels, kernel_size=1)
        self._finish_init(out_channels, activation)

    def forward(self, X, prompt_heatmap):
        if self.training:
            raise RuntimeError("unet-b200 is inference-only (BatchNorm is folded): call model.eval() first")
        if not X.is_cuda or not prompt_heatmap.is_cuda:
            raise RuntimeError("unet-b200 has no CPU path: inputs must be CUDA tensors")
        with torch.no_grad():
            clip_features = self.clip_feature_extractor(X)
            logits = self.engine(X.device).forward_prompt(X, prompt_heatmap, clip_features)
        return self.activation(logits)
