"""Drop-in for the reference's models/autoencoder.py `Encoder` (:5-26): the UNet encoder half as its own
module, so that checkpoints keyed `encoder.input.*`, `encoder.enc1.*`, ... load unchanged.

Inside `pre_trained.SegmentationModel.forward` the encoder and the decoder run as ONE pass of the sm_100a library
(the same launches as `UNet`, feature maps never leave the device pipeline).  On its own, `Encoder.forward(x)` returns
the reference's dict of feature maps {x0, enc1, enc2, enc3, bottleneck} (fp32 NCHW) through `ub_encoder_forward`
(`UB_ARCH_ENCODER`: stem, enc1-3, bottleneck on the tensor-core kernel, then an un-block kernel per map), so callers
that take encoder features (the reference's pre-training scripts) can switch imports.

`Decoder` (:28-42) and `Autoencoder` (:45-58) — the reconstruction model the encoder is pre-trained with: three
`ConvBlockUpsample` stages without skips, a 1x1 conv 64 -> out_channels and a sigmoid.  `Autoencoder.forward` runs the
whole pass in the library (`UB_ARCH_AUTOENCODER`, `ub_forward`): stem, enc1-3, bottleneck, three ConvT + double-conv
stages on the tensor-core kernel, then `out1x1_sigmoid_kernel`.  Inference only.
"""
from __future__ import annotations

import torch.nn as nn

import torch

from .processing_blocks import ConvBlock, ConvBlockDownsample, ConvBlockUpsample, _no_standalone
from .UNet import _B200Net


class Encoder(_B200Net):
    """input 1x1 (3->32), enc1..3 (ConvBlock + 2x2 max-pool), bottleneck 256->512; reference :5-26.
    forward(x) -> {"x0": [B,32,H,W], "enc1": [B,64,H/2,W/2], "enc2": [B,128,H/4,W/4], "enc3": [B,256,H/8,W/8],
    "bottleneck": [B,512,H/8,W/8]} (:17-25)."""
    _arch = "encoder"

    def __init__(self, in_channels=3):
        super().__init__()
        self.input = nn.Conv2d(in_channels, 32, kernel_size=1, padding=0)
        self.enc1 = ConvBlockDownsample(32, 64)
        self.enc2 = ConvBlockDownsample(64, 128)
        self.enc3 = ConvBlockDownsample(128, 256)
        self.bottleneck = ConvBlock(256, 512)
        self._n_classes = 3      # (unused by this half; the context wants a class count)
        self._engines = {}

    def forward(self, x):
        if self.training:
            raise RuntimeError("unet-b200 is inference-only (BatchNorm is folded): call model.eval() first")
        if not x.is_cuda:
            raise RuntimeError("unet-b200 has no CPU path: inputs must be CUDA tensors")
        with torch.no_grad():
            return self.engine(x.device).encoder_forward(x)


class Decoder(nn.Module):
    """dec1..3 (ConvT 2x2 -> ConvBlock, no skip) + 1x1 conv 64 -> out_channels; reference :28-42.  Container only."""

    def __init__(self, out_channels=3):
        super().__init__()
        self.dec1 = ConvBlockUpsample(512, 256)
        self.dec2 = ConvBlockUpsample(256, 128)
        self.dec3 = ConvBlockUpsample(128, 64)
        self.out = nn.Conv2d(64, out_channels, kernel_size=1, padding=0)

    def forward(self, bottleneck):
        _no_standalone("Decoder")


class Autoencoder(_B200Net):
    """`Autoencoder(in_channels, out_channels)(x) -> sigmoid(reconstruction)` [B,out_channels,H,W]; reference :45-58."""
    _arch = "autoencoder"

    def __init__(self, in_channels=3, out_channels=3):
        super().__init__()
        self.encoder = Encoder(in_channels)
        self.decoder = Decoder(out_channels)
        self._finish_init(out_channels, nn.Identity())  # the sigmoid is applied inside the library

    @staticmethod
    def _flat_name(key: str) -> str:
        for prefix in ("encoder.", "decoder."):
            if key.startswith(prefix):
                return key[len(prefix):]
        return key
