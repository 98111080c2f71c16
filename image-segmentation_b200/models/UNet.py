"""Drop-in for the reference's models/UNet.py: `UNet` (:7-76) and `LargeUNet` (:78-148).

Same constructor, same sub-module tree (hence the same `state_dict` keys, shapes and default
initialisation order), same `forward(x[B,3,H,W] fp32) -> logits[B,C,H,W] fp32`; the computation is the
hand-written sm_100a pipeline behind include/unet_b200.h (BatchNorm folded at load, bf16 tensor-core
convolutions with fp32 accumulation).  Inference only, CUDA sm_100a only.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import runtime
from .processing_blocks import ConvBlock, ConvBlockDownsample, ConvBlockUpsampleSkip


class _B200Net(nn.Module):
    _arch = "unet"
    _host_side_prefixes = ()  # sub-modules that compute outside the library (their weights are not handed to it)

    def _finish_init(self, out_channels, activation):
        self.activation = activation
        self._n_classes = out_channels
        self._engines = {}  # (device, max_microbatch) -> (weights key, Engine)

    def _weights_key(self):
        return tuple((p.data_ptr(), p._version) for p in list(self.parameters()) + list(self.buffers()))

    def engine(self, device=None, max_microbatch=None):
        """The library context holding this module's folded+packed weights (rebuilt when they change)."""
        dev = torch.device(device) if device is not None else next(self.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("unet-b200 has no CPU path: move the model and inputs to a CUDA (sm_100a) device")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        mb = max_microbatch or runtime.default_microbatch()
        wkey = self._weights_key()
        slot = self._engines.get((dev, mb))
        # one context per (device, micro-batch): asking for another micro-batch (e.g. a plain model(x) beside a
        # RobustnessEvaluator that holds its own engine) never destroys a context somebody else is using; only a
        # change of the WEIGHTS retires a context (its holders then get Engine's "rebuilt" error, not a stale result)
        if slot is None or slot[0] != wkey:
            if slot is not None:
                slot[1].close("the model's weights changed and its engine was rebuilt: call model.engine() again")
            eng = runtime.new_engine(dev, self._arch, 3, self._n_classes, mb)
            eng.load_state_dict({self._flat_name(k): v for k, v in self.state_dict().items()
                                 if not k.startswith(self._host_side_prefixes)})
            eng.set_layer_couts(self.layer_couts())
            self._engines[(dev, mb)] = (wkey, eng)
            slot = self._engines[(dev, mb)]
        return slot[1]

    @staticmethod
    def _flat_name(key: str) -> str:
        """state_dict key -> the library's layer name (the UNet key); identity for UNet/LargeUNet."""
        return key

    def layer_couts(self):
        out = {}
        for name, m in self.named_modules():
            if name.startswith(self._host_side_prefixes):
                continue
            name = self._flat_name(name)
            if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)) and name not in ("input", "out"):
                out[name] = m.out_channels
        return out

    def forward(self, x):
        if self.training:
            raise RuntimeError("unet-b200 is inference-only (BatchNorm is folded): call model.eval() first")
        if not x.is_cuda:
            raise RuntimeError("unet-b200 has no CPU path: inputs must be CUDA tensors")
        eng = self.engine(x.device)
        with torch.no_grad():
            logits = eng.forward(x)
        return self.activation(logits)


class UNet(_B200Net):
    """32-64-128-256 encoder, 512 bottleneck; reference models/UNet.py:37-76."""
    _arch = "unet"

    def __init__(self, in_channels=3, out_channels=3, activation=nn.Identity()):
        super().__init__()
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


class LargeUNet(_B200Net):
    """One more stage (… 512, bottleneck 1024); reference models/UNet.py:103-148."""
    _arch = "large"

    def __init__(self, in_channels=3, out_channels=3, activation=nn.Identity()):
        super().__init__()
        self.input = nn.Conv2d(in_channels, 32, kernel_size=1, padding=0)
        self.enc1 = ConvBlockDownsample(32, 64)
        self.enc2 = ConvBlockDownsample(64, 128)
        self.enc3 = ConvBlockDownsample(128, 256)
        self.enc4 = ConvBlockDownsample(256, 512)
        self.bottleneck = ConvBlock(512, 1024)
        self.dec1 = ConvBlockUpsampleSkip(1024, 512)
        self.dec2 = ConvBlockUpsampleSkip(512, 256)
        self.dec3 = ConvBlockUpsampleSkip(256, 128)
        self.dec4 = ConvBlockUpsampleSkip(128, 64)
        self.dec5 = ConvBlockUpsampleSkip(64, 32)
        self.out = nn.Conv2d(32, out_channels, kernel_size=1, padding=0)
        self._finish_init(out_channels, activation)
