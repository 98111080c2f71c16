"""Drop-in for the reference's models/CLIP_models.py `ClipUnet` (:65-134) — the model its
scripts/robustness_evaluation.py evaluates (:36-42).

Same constructor, same sub-module tree and state_dict keys.  The forward pass is
    clip_features = CLIP ViT-B/32 image embedding of X            (HuggingFace transformers, as in the reference)
    logits        = unet-b200 library: stem, enc1-3, cross-attention fusion of clip_features, dec1-4, head
The reference also runs `self.bottleneck(enc3)` (:122) but never uses its value: `CrossAttentionFusion` attends from
the bottleneck positions to a single CLIP token, so its output does not depend on the queries
(models/processing_blocks.py:310-322).  The `bottleneck.*` parameters are kept (checkpoints load) and not computed.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .processing_blocks import (ClipFeatureExtractor, ConvBlock, ConvBlockDownsample, ConvBlockUpsampleSkip,
                                CrossAttentionFusion)
from .UNet import _B200Net


class ClipUnet(_B200Net):
    _arch = "clip"
    _host_side_prefixes = ("clip_feature_extractor.",)

    def __init__(self, out_channels=3, in_channels=3, activation=nn.Identity(), clip_model=None, clip_autocast_dtype=None, clip_backend=None):
        super().__init__()
        self.clip_feature_extractor = ClipFeatureExtractor(train=False, clip_model=clip_model,
                                                           autocast_dtype=clip_autocast_dtype, backend=clip_backend)
        self.cross_attention_fusion = CrossAttentionFusion(512, num_heads=1)
        self.input = nn.Conv2d(in_channels, 32, kernel_size=1, padding=0)
        self.enc1 = ConvBlockDownsample(32, 64)
        self.enc2 = ConvBlockDownsample(64, 128)
        self.enc3 = ConvBlockDownsample(128, 256)
        self.bottleneck = ConvBlock(256, 512)
        self.dec1 = ConvBlockUpsampleSkip(512, 256)
        self.dec2 = ConvBlockUpsampleSkip(256, 128)
        self.dec3 = ConvBlockUpsampleSkip(128, 64)
        self.dec4 = ConvBlockUpsampleSkip(64, 32)
        self.out = nn.Conv2d(32, out_channels, kernel_size=1, padding=0)
        self._finish_init(out_channels, activation)

    def _weights_key(self):
        # the CLIP encoder's weights never reach the library: only the rest decides when it is re-packed
        named = [(n, t) for n, t in list(self.named_parameters()) + list(self.named_buffers())
                 if not n.startswith(self._host_side_prefixes)]
        return tuple((t.data_ptr(), t._version) for _, t in named)

    def forward(self, X):
        if self.training:
            raise RuntimeError("unet-b200 is inference-only (BatchNorm is folded): call model.eval() first")
        if not X.is_cuda:
            raise RuntimeError("unet-b200 has no CPU path: inputs must be CUDA tensors")
        with torch.no_grad():
            clip_features = self.clip_feature_extractor(X)
            logits = self.engine(X.device).forward(X, clip_feats=clip_features)
        return self.activation(logits)
