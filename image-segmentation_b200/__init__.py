"""unet-b200: B200-native (sm_100a) robustness-evaluation path of MattiDeBeer/image-segmentation.

Sub-modules mirror the reference's layout so a user can switch imports one for one:
  models.UNet, models.processing_blocks, models.losses, models.model_wrappers,
  customDatasets.perturbations, customDatasets.datasets, scripts.robustness_evaluation.
All compute goes through the C-ABI library csrc/libunetb200.so (see include/unet_b200.h); there is
no CPU or PyTorch fallback for the network, the perturbations or the metrics on that path.
"""
__version__ = "0.1.0"
