"""Drop-in for the reference's models/pre_trained.py: `SegmentationDecoderSkip` (:7-24) and
`SegmentationModel` (:47-56) — a (pre-trained, frozen) `autoencoder.Encoder` plus a skip decoder.

The arithmetic is the `UNet`'s (same layers, same shapes, same order); only the parameter tree differs:
`encoder.<name>` / `decoder.<name>` instead of `<name>`.  `SegmentationModel` therefore hands the library
its state_dict with those prefixes dropped and runs the identical fused sm_100a pass; the dict of
encoder features the reference passes between the halves stays in device buffers.  The halves also work on their own
(`Encoder.forward` -> dict, `SegmentationDecoderSkip.forward(dict)` -> logits) through `ub_encoder_forward` /
`ub_skip_decoder_forward`.
"""
from __future__ import annotations

import torch.nn as nn

from .processing_blocks import ConvBlockUpsampleSkip, _no_standalone
from .UNet import _B200Net


class SegmentationDecoderSkip(_B200Net):
    """dec1..4 (ConvT 2x2 -> resize to skip -> concat -> ConvBlock) + 1x1 head; reference :7-24.
    forward(features) takes `Encoder.forward`'s dict and returns the logits (`ub_skip_decoder_forward`)."""
    _arch = "skip_decoder"

    def __init__(self, out_channels):
        super().__init__()
        self.dec1 = ConvBlockUpsampleSkip(512, 256)
        self.dec2 = ConvBlockUpsampleSkip(256, 128)
        self.dec3 = ConvBlockUpsampleSkip(128, 64)
        self.dec4 = ConvBlockUpsampleSkip(64, 32)
        self.out = nn.Conv2d(32, out_channels, kernel_size=1, padding=0)
        self._n_classes = out_channels
        self._engines = {}

    def forward(self, features):
        if self.training:
            raise RuntimeError("unet-b200 is inference-only (BatchNorm is folded): call model.eval() first")
        x0 = features["x0"]
        if not x0.is_cuda:
            raise RuntimeError("unet-b200 has no CPU path: inputs must be CUDA tensors")
        import torch
        with torch.no_grad():
            return self.engine(x0.device).skip_decoder_forward(features)


class SegmentationModel(_B200Net):
    """`SegmentationModel(encoder, decoder)(x) -> logits`; reference :47-56 (no output activation)."""
    _arch = "unet"

    def __init__(self, encoder, decoder):
        super().__init__()
        if not isinstance(decoder, SegmentationDecoderSkip):
            raise TypeError("unet-b200 implements SegmentationModel with a SegmentationDecoderSkip decoder only")
        self.encoder = encoder
        self.decoder = decoder
        self._finish_init(decoder.out.out_channels, nn.Identity())

    @staticmethod
    def _flat_name(key: str) -> str:
        for prefix in ("encoder.", "decoder."):
            if key.startswith(prefix):
                return key[len(prefix):]
        return key
