"""Generates tests/golden/*.npz by RUNNING THE REFERENCE'S OWN CODE (imported from /root/reference
through oracle/refload.py).  Run in the build container only:  python -m oracle.make_golden

  perturb_u8.npz   customDatasets/perturbations.py wrappers, all 8 families x the 10 script levels
                   (scripts/robustness_evaluation.py:59-92) on 2 synthetic 3x24x40 images, with the RNG
                   artefacts (noise tensors, occlusion corners, salt-pepper code maps) captured by
                   replaying random.seed(42)/torch.manual_seed(42) (robustness_evaluation.py:32-34).
  perturb_f.npz    models/processing_blocks.py float-space modules (all but RepeatedBlur, which needs
                   kornia) on a [2,3,24,40] batch, EVERY level TestWrapper.test_robustness visits
                   (models/model_wrappers.py:740-764: 69 of its 79 levels; the other 10 are RepeatedBlur).
  unet_small.npz   reference UNet (seed 42 + randomised BatchNorm) logits on a [2,3,32,64] input, the
                   reference IoU / PixelAccuracy on them, and a checksum of the seeded state_dict.
  dataset_rows.npz reference CustomImageDataset._deserialize_datapoint on 6 synthetic rows (trimap remap), masks +
                   a checksum of the decoded fp32 images.
  autoencoder_small.npz  reference Autoencoder (models/autoencoder.py:45-58): keys, checksums, sigmoid reconstruction of [2,3,32,64]
  clipunet_prompt_small.npz  reference ClipUnetPrompt (same stand-in encoder) on 2 images + Gaussian heat-maps: keys, embedding, logits
  clipunet_small.npz  reference ClipUnet with a seeded random-init CLIP encoder (the checkpoint needs the network): key
                   list, CLIP embedding and logits on the [2,3,32,64] input.
  segmodel_small.npz  reference SegmentationModel(Encoder(), SegmentationDecoderSkip(3)) (models/pre_trained.py:47-56,
                   models/autoencoder.py:5-26), seed 42 + randomised BatchNorm: state_dict key list, checksums and
                   logits on the same [2,3,32,64] input.
"""
import hashlib
import os
import random

import numpy as np
import torch

from oracle import refload

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

SCRIPT_LEVELS = {  # scripts/robustness_evaluation.py:59-92
    "gaussian_noise": ("GaussianPixelNoise", [0, 2, 4, 6, 8, 10, 12, 14, 16, 18]),
    "gaussian_blur": ("GaussianBlur", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9]),
    "contrast_increase": ("ContrastIncrease", [1.0, 1.01, 1.02, 1.03, 1.04, 1.05, 1.1, 1.15, 1.2, 1.25]),
    "contrast_decrease": ("ContrastDecrease", [1.0, 0.95, 0.9, 0.85, 0.8, 0.6, 0.4, 0.3, 0.2, 0.1]),
    "brightness_increase": ("BrightnessIncrease", [0, 5, 10, 15, 20, 25, 30, 35, 40, 45]),
    "brightness_decrease": ("BrightnessDecrease", [0, 5, 10, 15, 20, 25, 30, 35, 40, 45]),
    "occlusion": ("OcclusionIncrease", [0, 5, 10, 15, 20, 25, 30, 35, 40, 45]),
    "salt_pepper_noise": ("SaltPepperNoise", [0.00, 0.02, 0.04, 0.06, 0.08, 0.10, 0.12, 0.14, 0.16, 0.18]),
}


WRAPPER_LEVELS = {  # models/model_wrappers.py:740-764 (class name -> every parameter value test_robustness passes)
    "GaussianPixelNoise": [1e-6, 2, 4, 6, 8, 10, 12, 14, 16, 18],
    "ContrastChange": [1.0, 1.01, 1.02, 1.03, 1.04, 1.05, 1.1, 1.15, 1.20, 1.25,
                       1.0, 0.95, 0.90, 0.85, 0.80, 0.60, 0.40, 0.30, 0.20, 0.10],
    "BrightnessChange": [0, 5, 10, 15, 20, 25, 30, 35, 40, 45, 0, -5, -10, -15, -20, -25, -30, -35, -40, -45],
    "Occlusion": [0, 5, 10, 15, 20, 25, 30, 35, 40, 45],
    "SaltAndPepper": [0.00, 0.02, 0.04, 0.06, 0.08, 0.10, 0.12, 0.14, 0.16],
}


class _Base(torch.utils.data.Dataset):
    def __init__(self, imgs):
        self.imgs = imgs

    def __len__(self):
        return len(self.imgs)

    def __getitem__(self, i):
        return self.imgs[i], torch.zeros(self.imgs[i].shape[1:], dtype=torch.long)


def small_images(n=2, h=24, w=40, seed=11):
    g = torch.Generator().manual_seed(seed)
    u8 = torch.randint(0, 256, (n, 3, h, w), generator=g, dtype=torch.uint8)
    return u8, u8.float() / 255.0


def state_checksum(sd):
    hsh = hashlib.sha256()
    for k in sorted(sd):
        if sd[k].is_floating_point():
            hsh.update(k.encode())
            hsh.update(sd[k].detach().cpu().float().numpy().tobytes())
    return hsh.hexdigest()


def golden_perturb_u8(ns):
    from oracle import perturb_oracle as po
    u8, imgs = small_images()
    out = {"images_u8": u8.numpy()}
    _, h, w = imgs[0].shape
    for fam, (cls_name, levels) in SCRIPT_LEVELS.items():
        cls = getattr(ns.perturbations, cls_name)
        for li, level in enumerate(levels):
            # reference run
            random.seed(42); torch.manual_seed(42)
            ds = cls(_Base(imgs), level)
            ref = torch.stack([ds[i][0] for i in range(len(ds))])
            ref_u8 = (ref * 255.0).round().to(torch.uint8)
            assert torch.equal(ref_u8.float() / 255.0, ref), (fam, level)  # outputs are exactly k/255
            out[f"{fam}/{li}/out"] = ref_u8.numpy()
            # RNG artefacts, replayed in the same order
            random.seed(42); torch.manual_seed(42)
            if fam == "gaussian_noise":
                out[f"{fam}/{li}/aux"] = np.stack([
                    torch.normal(mean=0.0, std=float(level), size=imgs[0].shape, dtype=torch.float).numpy()
                    for _ in range(len(imgs))])
            elif fam == "occlusion":
                out[f"{fam}/{li}/aux"] = np.array([po.replay_occlusion_corner(level, h, w) for _ in range(len(imgs))],
                                                 dtype=np.int32)
            elif fam == "salt_pepper_noise":
                out[f"{fam}/{li}/aux"] = np.stack([po.replay_salt_pepper_code(level, h, w) for _ in range(len(imgs))])
    np.savez_compressed(os.path.join(OUT, "perturb_u8.npz"), **out)
    return out


def golden_perturb_f(ns):
    from oracle import perturb_oracle as po
    u8, imgs = small_images(seed=12)
    out = {"images_u8": u8.numpy()}
    b, _, h, w = imgs.shape
    for fam, levels in WRAPPER_LEVELS.items():
        for li, level in enumerate(levels):
            random.seed(42); torch.manual_seed(42)
            mod = getattr(ns.blocks, fam)(level)
            ref = mod(imgs.clone())
            out[f"{fam}/{li}/out"] = ref.numpy()
            random.seed(42); torch.manual_seed(42)
            if fam == "GaussianPixelNoise":
                out[f"{fam}/{li}/aux"] = torch.randn_like(imgs).numpy()
            elif fam == "Occlusion":
                out[f"{fam}/{li}/aux"] = np.array(po.replay_f_occlusion_corners(b, level, h, w), dtype=np.int32)
            elif fam == "SaltAndPepper":
                out[f"{fam}/{li}/aux"] = torch.rand((b, 1, h, w)).numpy()
            out[f"{fam}/{li}/level"] = np.float64(level)
    np.savez_compressed(os.path.join(OUT, "perturb_f.npz"), **out)
    return out


def golden_unet(ns):
    from image_segmentation_b200 import synthetic
    torch.manual_seed(42)
    m = ns.UNet().eval()
    base_sum = state_checksum(m.state_dict())
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    m.load_state_dict(sd)
    imgs, masks = synthetic.make_dataset(2, 32, 64, seed=5)
    x = synthetic.to_float_nchw(imgs)
    with torch.no_grad():
        logits = m(x)
        # centre the head so that all three classes occur, then re-run
        sd["out.bias"] = sd["out.bias"] - logits.mean(dim=(0, 2, 3))
        m.load_state_dict(sd)
        logits = m(x)
        iou = ns.losses.IoU()(logits, masks.long())
        pa = ns.losses.PixelAccuracy()(logits, masks.long())
    out = dict(images_u8=imgs.numpy(), masks_u8=masks.numpy(), logits=logits.numpy(), iou=iou.numpy(),
               pixel_accuracy=pa.numpy(), out_bias=sd["out.bias"].numpy(),
               seed42_state_sha256=np.array(base_sum), final_state_sha256=np.array(state_checksum(sd)))
    np.savez_compressed(os.path.join(OUT, "unet_small.npz"), **out)
    return out


def golden_segmodel(ns):
    from image_segmentation_b200 import synthetic
    torch.manual_seed(42)
    m = ns.pre_trained.SegmentationModel(ns.autoencoder.Encoder(), ns.pre_trained.SegmentationDecoderSkip(3)).eval()
    base_sum = state_checksum(m.state_dict())
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    m.load_state_dict(sd)
    imgs, masks = synthetic.make_dataset(2, 32, 64, seed=5)
    x = synthetic.to_float_nchw(imgs)
    with torch.no_grad():
        logits = m(x)
        sd["decoder.out.bias"] = sd["decoder.out.bias"] - logits.mean(dim=(0, 2, 3))
        m.load_state_dict(sd)
        logits = m(x)
    keys = list(m.state_dict().keys())
    out = dict(images_u8=imgs.numpy(), logits=logits.numpy(), out_bias=sd["decoder.out.bias"].numpy(),
               keys=np.array(keys), shapes=np.array([str(tuple(m.state_dict()[k].shape)) for k in keys]),
               seed42_state_sha256=np.array(base_sum), final_state_sha256=np.array(state_checksum(sd)))
    np.savez_compressed(os.path.join(OUT, "segmodel_small.npz"), **out)
    return out


def golden_dataset_rows(ns):
    """The reference's own `_deserialize_datapoint` (customDatasets/datasets.py:92-131), called unbound on 6
    synthetic rows (the HF download its __init__ does is not possible here; the method only uses
    `self._deserialize_numpy`)."""
    from image_segmentation_b200 import synthetic
    ds_mod = __import__("customDatasets.datasets", fromlist=["CustomImageDataset"])
    cls = ds_mod.CustomImageDataset
    obj = cls.__new__(cls)
    rows = synthetic.make_raw_rows(6, 256, 256, seed=21)
    masks, img_sha = [], hashlib.sha256()
    for r in rows:
        image, mask = cls._deserialize_datapoint(obj, r)
        assert image.dtype == torch.float32 and tuple(image.shape) == (3, 256, 256)
        img_sha.update(image.numpy().tobytes())
        masks.append(mask.numpy().astype(np.uint8))
    out = dict(masks=np.stack(masks), images_sha256=np.array(img_sha.hexdigest()),
               raw_sha256=np.array(hashlib.sha256(b"".join(r["image"] + r["mask"] for r in rows)).hexdigest()))
    np.savez_compressed(os.path.join(OUT, "dataset_rows.npz"), **out)
    return out


def golden_clipunet():
    """Reference ClipUnet (models/CLIP_models.py:65-134) with a seeded random-init CLIP ViT-B/32 (refload.load_clip_unet):
    key list of the non-CLIP part, the CLIP embedding of the input and the logits."""
    from image_segmentation_b200 import synthetic
    ns = refload.load_clip_unet(clip_seed=0)
    torch.manual_seed(42)
    m = ns.ClipUnet().to("cpu").eval()
    own = {k: v for k, v in m.state_dict().items() if not k.startswith("clip_feature_extractor.")}
    base_sum = state_checksum(own)
    sd = synthetic.randomize_batchnorm(own, seed=7)
    m.load_state_dict(sd, strict=False)
    imgs, _ = synthetic.make_dataset(2, 32, 64, seed=5)
    x = synthetic.to_float_nchw(imgs)
    with torch.no_grad():
        feats = m.clip_feature_extractor(x.to(m.clip_feature_extractor.device)).cpu()
        logits = m(x)
        sd["out.bias"] = sd["out.bias"] - logits.mean(dim=(0, 2, 3))
        m.load_state_dict(sd, strict=False)
        logits = m(x)
    keys = list(own.keys())
    out = dict(images_u8=imgs.numpy(), clip_feats=feats.numpy(), logits=logits.numpy(), out_bias=sd["out.bias"].numpy(),
               keys=np.array(keys), shapes=np.array([str(tuple(own[k].shape)) for k in keys]),
               seed42_state_sha256=np.array(base_sum), final_state_sha256=np.array(state_checksum(sd)))
    np.savez_compressed(os.path.join(OUT, "clipunet_small.npz"), **out)
    return out


def golden_autoencoder(ns):
    """Reference Autoencoder (models/autoencoder.py:45-58), seed 42 + randomised BatchNorm: key list, checksums and the
    sigmoid reconstruction of the [2,3,32,64] input."""
    from image_segmentation_b200 import synthetic
    torch.manual_seed(42)
    m = ns.autoencoder.Autoencoder().eval()
    base = {k: v.clone() for k, v in m.state_dict().items()}
    sd = synthetic.randomize_batchnorm(base, seed=7)
    # default init leaves the pre-sigmoid output within +-0.05 (everything ~0.5): widen it so the test sees structure
    sd["decoder.out.weight"] = sd["decoder.out.weight"] * 8.0
    m.load_state_dict(sd)
    imgs, _ = synthetic.make_dataset(2, 32, 64, seed=5)
    x = synthetic.to_float_nchw(imgs)
    with torch.no_grad():
        recon = m(x)
    keys = list(base.keys())
    out = dict(images_u8=imgs.numpy(), recon=recon.numpy(), keys=np.array(keys),
               shapes=np.array([str(tuple(base[k].shape)) for k in keys]),
               seed42_state_sha256=np.array(state_checksum(base)), final_state_sha256=np.array(state_checksum(sd)))
    np.savez_compressed(os.path.join(OUT, "autoencoder_small.npz"), **out)
    return out


def golden_clipunet_prompt():
    """Reference ClipUnetPrompt (models/prompt_segmentation.py:31-95) with the seeded stand-in CLIP encoder: key list of
    the non-CLIP part, the CLIP embedding, the Gaussian prompt heat-maps (datasets.py:599-610 form) and the logits."""
    from image_segmentation_b200 import synthetic
    ns = refload.load_clip_unet(clip_seed=0)
    torch.manual_seed(42)
    m = ns.ClipUnetPrompt().to("cpu").eval()
    own = {k: v for k, v in m.state_dict().items() if not k.startswith("clip_feature_extractor.")}
    base_sum = state_checksum(own)
    # amplified so that the prompt and the CLIP embedding visibly move the logits (see synthetic.amplify_bottleneck_path)
    sd = synthetic.amplify_bottleneck_path(synthetic.randomize_batchnorm(own, seed=7))
    m.load_state_dict(sd, strict=False)
    imgs, _ = synthetic.make_dataset(2, 32, 64, seed=5)
    x = synthetic.to_float_nchw(imgs)
    heat = synthetic.prompt_heatmaps(2, 32, 64, seed=3, sigma=3.0)
    with torch.no_grad():
        feats = m.clip_feature_extractor(x.to(m.clip_feature_extractor.device)).cpu()
        logits = m(x, heat)
        logits_flipped = m(x, heat.flip(-1).contiguous())   # another click position: the prompt branch matters
    keys = list(own.keys())
    out = dict(images_u8=imgs.numpy(), heatmaps=heat.numpy(), clip_feats=feats.numpy(), logits=logits.numpy(),
               logits_flipped_prompt=logits_flipped.numpy(),
               keys=np.array(keys), shapes=np.array([str(tuple(own[k].shape)) for k in keys]),
               seed42_state_sha256=np.array(base_sum), final_state_sha256=np.array(state_checksum(sd)))
    np.savez_compressed(os.path.join(OUT, "clipunet_prompt_small.npz"), **out)
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    ns = refload.load()
    golden_perturb_u8(ns)
    golden_perturb_f(ns)
    golden_unet(ns)
    golden_segmodel(ns)
    golden_autoencoder(ns)
    golden_dataset_rows(ns)
    golden_clipunet()
    golden_clipunet_prompt()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
