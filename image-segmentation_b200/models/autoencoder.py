"""Drop-in for the reference's models/autoencoder.py `Encoder` (:5-26): the UNet encoder half as its own
module, so that checkpoints keyed `encoder.input.*`, `encoder.enc1.*`, ... load unchanged.

In unet-b200 the encoder is a PARAMETER CONTAINER: it computes as part of
`pre_trained.SegmentationModel.forward`, which runs encoder and decoder in one pass of the sm_100a library
(the same launches as `UNet`).  The stand-alone `Encoder.forward` of the reference returns the dict of
feature maps {x0, enc1, enc2, enc3, bottleneck}; those never leave the device pipeline here, so calling it
on its own raises.  The reconstruction `Decoder`/`Autoencoder` (:28-58, sigmoid output, no skips) is a
training-time model outside the robustness-evaluation path and is not provided.
"""
from __future__ import annotations

import torch.nn as nn

from .processing_blocks import ConvBlock, ConvBlockDownsample, _no_standalone


class Encoder(nn.Module):
    """input 1x1 (3->32), enc1..3 (ConvBlock + 2x2 max-pool), bottleneck 256->512; reference :5-26."""

    def __init__(self, in_channels=3):
        super().__init__()
        self.input = nn.Conv2d(in_channels, 32, kernel_size=1, padding=0)
        self.enc1 = ConvBlockDownsample(32, 64)
        self.enc2 = ConvBlockDownsample(64, 128)
        self.enc3 = ConvBlockDownsample(128, 256)
        self.bottleneck = ConvBlock(256, 512)

    def forward(self, x):
        _no_standalone("Encoder")
