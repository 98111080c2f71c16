"""Pins the oracle: (a) against the golden vectors produced by the reference's own code
(tests/golden/*.npz, made by oracle/make_golden.py), always; (b) against the live reference modules
when /root/reference is present (build container only)."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import metrics_oracle as mo
from oracle import perturb_oracle as po
from oracle import refload, unet_oracle
from oracle.make_golden import SCRIPT_LEVELS, state_checksum

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def g_u8():
    return np.load(os.path.join(GOLD, "perturb_u8.npz"))


@pytest.fixture(scope="module")
def g_f():
    return np.load(os.path.join(GOLD, "perturb_f.npz"))


def oracle_u8(fam, level, u8, aux):
    if fam == "gaussian_noise":
        return po.u8_gaussian_noise(u8, aux)
    if fam == "gaussian_blur":
        return po.u8_blur(u8, level)
    if fam in ("contrast_increase", "contrast_decrease"):
        return po.u8_contrast(u8, level)
    if fam == "brightness_increase":
        return po.u8_brightness(u8, level, True)
    if fam == "brightness_decrease":
        return po.u8_brightness(u8, level, False)
    if fam == "occlusion":
        return np.stack([po.u8_occlusion(u8[i], level, int(aux[i][0]), int(aux[i][1])) for i in range(len(u8))])
    if fam == "salt_pepper_noise":
        return np.stack([po.u8_salt_pepper(u8[i], aux[i]) for i in range(len(u8))])
    raise KeyError(fam)


@pytest.mark.parametrize("fam", list(SCRIPT_LEVELS))
def test_u8_perturbations_match_golden(g_u8, fam):
    u8 = g_u8["images_u8"]
    for li, level in enumerate(SCRIPT_LEVELS[fam][1]):
        aux = g_u8[f"{fam}/{li}/aux"] if f"{fam}/{li}/aux" in g_u8.files else None
        got = oracle_u8(fam, level, u8, aux)
        assert np.array_equal(got, g_u8[f"{fam}/{li}/out"]), (fam, level)


def oracle_f(fam, level, x, aux):
    if fam == "GaussianPixelNoise":
        return po.f_gaussian_noise(x, aux, level)
    if fam == "ContrastChange":
        return po.f_contrast(x, level)
    if fam == "BrightnessChange":
        return po.f_brightness(x, level)
    if fam == "Occlusion":
        return po.f_occlusion(x, int(level), [tuple(int(v) for v in c) for c in aux])
    if fam == "SaltAndPepper":
        return po.f_salt_pepper(x, aux, level)
    raise KeyError(fam)


def _float_levels():
    from oracle.make_golden import WRAPPER_LEVELS
    return [(fam, li) for fam, levels in WRAPPER_LEVELS.items() for li in range(len(levels))]


@pytest.mark.parametrize("fam,li", _float_levels())
def test_float_perturbations_match_golden(g_f, fam, li):
    """Every level TestWrapper.test_robustness visits (models/model_wrappers.py:740-764) except RepeatedBlur's."""
    from oracle.make_golden import WRAPPER_LEVELS
    x = g_f["images_u8"].astype(np.float32) / np.float32(255.0)
    level = float(g_f[f"{fam}/{li}/level"])
    assert level == float(WRAPPER_LEVELS[fam][li])
    aux = g_f[f"{fam}/{li}/aux"] if f"{fam}/{li}/aux" in g_f.files else None
    got = oracle_f(fam, level, x, aux)
    assert np.array_equal(got, g_f[f"{fam}/{li}/out"]), (fam, level)  # bit-exact fp32


def test_unet_and_metrics_match_golden():
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.UNet import UNet
    g = np.load(os.path.join(GOLD, "unet_small.npz"))
    torch.manual_seed(42)
    m = UNet()
    assert state_checksum(m.state_dict()) == str(g["seed42_state_sha256"])  # same init as the reference
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    sd["out.bias"] = torch.from_numpy(g["out_bias"])
    assert state_checksum(sd) == str(g["final_state_sha256"])
    imgs, masks = synthetic.make_dataset(2, 32, 64, seed=5)
    assert np.array_equal(imgs.numpy(), g["images_u8"]) and np.array_equal(masks.numpy(), g["masks_u8"])
    logits = unet_oracle.forward(sd, synthetic.to_float_nchw(imgs))
    ref = torch.from_numpy(g["logits"])
    assert float((logits - ref).abs().max()) <= 2e-5 * float(ref.abs().max())  # fp32, thread-count noise only
    pred = unet_oracle.hard_prediction(ref).numpy()
    cm = mo.confusion_matrix(pred, masks.numpy())
    assert mo.iou_from_cm(cm) == g["iou"][()]
    assert mo.pixel_accuracy_from_cm(cm) == g["pixel_accuracy"][()]
    assert float(mo.iou_direct(ref, masks.long())) == float(g["iou"])
    assert float(mo.pixel_accuracy_direct(ref, masks.long())) == float(g["pixel_accuracy"])


def test_segmentation_model_tree_and_oracle_against_reference_golden():
    """SegmentationModel(Encoder, SegmentationDecoderSkip) (reference models/pre_trained.py:47-56): same key
    list, shapes and seeded initialisation as the reference's, and the oracle network on the prefix-stripped
    state_dict reproduces the reference's logits."""
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.autoencoder import Encoder
    from image_segmentation_b200.models.pre_trained import SegmentationDecoderSkip, SegmentationModel
    from oracle.make_golden import state_checksum
    g = np.load(os.path.join(GOLD, "segmodel_small.npz"))
    torch.manual_seed(42)
    m = SegmentationModel(Encoder(), SegmentationDecoderSkip(3))
    sd0 = m.state_dict()
    assert list(sd0.keys()) == [str(k) for k in g["keys"]]
    assert [str(tuple(v.shape)) for v in sd0.values()] == [str(s) for s in g["shapes"]]
    assert state_checksum(sd0) == str(g["seed42_state_sha256"])
    sd = synthetic.randomize_batchnorm(sd0, seed=7)
    sd["decoder.out.bias"] = torch.from_numpy(g["out_bias"])
    assert state_checksum(sd) == str(g["final_state_sha256"])
    flat = {m._flat_name(k): v for k, v in sd.items()}
    logits = unet_oracle.forward(flat, synthetic.to_float_nchw(torch.from_numpy(g["images_u8"])))
    ref = torch.from_numpy(g["logits"])
    assert float((logits - ref).abs().max()) <= 2e-5 * float(ref.abs().max())
    with pytest.raises(RuntimeError):
        m.encoder(torch.zeros(1, 3, 32, 32))  # parameter container, no per-half fallback


def test_dataset_row_decoding_against_reference_golden():
    """oracle.dataset_oracle vs the reference's CustomImageDataset._deserialize_datapoint (datasets.py:107-131)."""
    import hashlib
    from image_segmentation_b200 import synthetic
    from oracle import dataset_oracle as do
    g = np.load(os.path.join(GOLD, "dataset_rows.npz"))
    rows = synthetic.make_raw_rows(6, 256, 256, seed=21)
    assert hashlib.sha256(b"".join(r["image"] + r["mask"] for r in rows)).hexdigest() == str(g["raw_sha256"])
    sha = hashlib.sha256()
    for i, r in enumerate(rows):
        image, mask = do.deserialize_datapoint(r)
        assert mask.dtype == np.int64 and np.array_equal(mask, g["masks"][i])
        sha.update(image.tobytes())
    assert sha.hexdigest() == str(g["images_sha256"])
    kinds = [set(np.unique(m)) for m in g["masks"]]
    assert kinds[0] == {0, 1} and kinds[1] == {0, 2} and kinds[2] == {0, 2}  # cat, dog, border-only row


def test_clip_fusion_restatement_equals_multihead_attention():
    """oracle.unet_oracle.clip_fusion vs nn.MultiheadAttention fed the way CrossAttentionFusion feeds it
    (processing_blocks.py:310-322): one key/value token per image -> the output ignores the queries."""
    torch.manual_seed(3)
    mha = torch.nn.MultiheadAttention(embed_dim=512, num_heads=1).eval()
    sd = {"cross_attention_fusion.cross_attn." + k: v for k, v in mha.state_dict().items()}
    feats = torch.randn(2, 512)
    q = torch.randn(6 * 5, 2, 512)
    with torch.no_grad():
        ref, _ = mha(q, feats.unsqueeze(0).expand(30, -1, -1), feats.unsqueeze(0).expand(30, -1, -1))
    ref = ref.permute(1, 2, 0).view(2, 512, 6, 5)
    got = unet_oracle.clip_fusion(sd, feats, 6, 5)
    assert float((got - ref).abs().max()) <= 1e-5 * float(ref.abs().max())


def test_clipunet_tree_and_oracle_against_reference_golden():
    """ClipUnet drop-in (reference models/CLIP_models.py:65-134): same keys/shapes/seeded init for everything but the
    CLIP encoder, same CLIP embedding from the same seeded encoder, and the oracle reproduces the reference's logits."""
    from transformers import CLIPConfig, CLIPModel
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.CLIP_models import ClipUnet
    from oracle.make_golden import state_checksum
    g = np.load(os.path.join(GOLD, "clipunet_small.npz"))
    torch.manual_seed(0)                      # oracle/refload.py: the stand-in checkpoint is CLIPConfig() at seed 0 ...
    clip = CLIPModel(CLIPConfig()).eval()
    m = ClipUnet(clip_model=clip).eval()      # ... and the rest of the model draws from the same stream, as in the reference
    own = {k: v for k, v in m.state_dict().items() if not k.startswith("clip_feature_extractor.")}
    assert list(own.keys()) == [str(k) for k in g["keys"]]
    assert [str(tuple(v.shape)) for v in own.values()] == [str(s) for s in g["shapes"]]
    assert state_checksum(own) == str(g["seed42_state_sha256"])
    sd = synthetic.randomize_batchnorm(own, seed=7)
    sd["out.bias"] = torch.from_numpy(g["out_bias"])
    assert state_checksum(sd) == str(g["final_state_sha256"])
    x = synthetic.to_float_nchw(torch.from_numpy(g["images_u8"]))
    if not torch.cuda.is_available():         # the extractor lives on the GPU when there is one
        from image_segmentation_b200.models.processing_blocks import ClipFeatureExtractor
        feats = ClipFeatureExtractor(clip_model=clip, backend="transformers")(x)   # (the default backend is GPU-only)
        assert float((feats - torch.from_numpy(g["clip_feats"])).abs().max()) <= 1e-4 * float(np.abs(g["clip_feats"]).max())
    logits = unet_oracle.forward(sd, x, clip_feats=torch.from_numpy(g["clip_feats"]))
    ref = torch.from_numpy(g["logits"])
    assert float((logits - ref).abs().max()) <= 2e-5 * float(ref.abs().max())
    with pytest.raises(RuntimeError):
        m(x)                                  # no CPU path


def test_clip_oracle_pinned_to_torchvision_and_the_reference_golden():
    """oracle/clip_oracle.py (the checker of ub_clip_image_features): its Resize + Normalize restatement equals
    torchvision's own transforms bit for bit (and the reference's CustomClipPreprocessor where /root/reference is
    present), its embedding of the golden images equals the embedding the REFERENCE's ClipFeatureExtractor produced
    (tests/golden/clipunet_small.npz, made by oracle/make_golden.py through the reference's class), and the tower
    written out from the state_dict equals transformers' own forward."""
    from torchvision import transforms
    from transformers import CLIPConfig, CLIPModel
    from image_segmentation_b200 import synthetic
    from oracle import clip_oracle as co
    gen = torch.Generator().manual_seed(3)
    tv = transforms.Compose([transforms.Resize((224, 224)), transforms.Normalize(mean=co.CLIP_MEAN, std=co.CLIP_STD)])
    for hw in ((256, 256), (224, 224), (200, 300), (97, 131), (512, 384)):
        x = torch.rand((2, 3) + hw, generator=gen)
        want = torch.stack([tv(img) for img in x])
        assert torch.equal(co.preprocess(x), want), hw
    if refload.available():
        ns = refload.load()
        ref_pre = ns.blocks.CustomClipPreprocessor(mean=list(co.CLIP_MEAN), std=list(co.CLIP_STD))
        x = torch.rand((2, 3, 256, 256), generator=gen)
        assert torch.equal(co.preprocess(x), ref_pre(x))
    g = np.load(os.path.join(GOLD, "clipunet_small.npz"))
    torch.manual_seed(0)
    clip = CLIPModel(CLIPConfig()).eval()     # the stand-in checkpoint of oracle/refload.py: CLIPConfig() at seed 0
    x = synthetic.to_float_nchw(torch.from_numpy(g["images_u8"]))
    feats = co.image_features(clip, x)
    ref = torch.from_numpy(g["clip_feats"])
    assert float((feats - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
    cfg = clip.config.vision_config
    restated = co.vit_restated(clip.state_dict(), co.preprocess(x), cfg.num_attention_heads, cfg.layer_norm_eps)
    assert float((restated - ref).abs().max()) <= 1e-4 * float(ref.abs().max())
    small = co.seeded_clip(seed=3, small=True)
    scfg = small.config.vision_config
    a = co.image_features(small, x)
    b = co.vit_restated(small.state_dict(), co.preprocess(x), scfg.num_attention_heads, scfg.layer_norm_eps)
    assert float((a - b).abs().max()) <= 1e-4 * float(a.abs().max())


def test_autoencoder_tree_and_oracle_against_reference_golden():
    """Autoencoder drop-in (reference models/autoencoder.py:28-58): same keys/shapes/seeded init, and the oracle's
    restatement reproduces the reference's sigmoid reconstruction."""
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.autoencoder import Autoencoder, Decoder
    from oracle.make_golden import state_checksum
    g = np.load(os.path.join(GOLD, "autoencoder_small.npz"))
    torch.manual_seed(42)
    m = Autoencoder().eval()
    own = m.state_dict()
    assert list(own.keys()) == [str(k) for k in g["keys"]]
    assert [str(tuple(v.shape)) for v in own.values()] == [str(s) for s in g["shapes"]]
    assert state_checksum(own) == str(g["seed42_state_sha256"])
    sd = synthetic.randomize_batchnorm(own, seed=7)
    sd["decoder.out.weight"] = sd["decoder.out.weight"] * 8.0
    assert state_checksum(sd) == str(g["final_state_sha256"])
    x = synthetic.to_float_nchw(torch.from_numpy(g["images_u8"]))
    recon = unet_oracle.autoencoder_forward(sd, x)
    assert float((recon - torch.from_numpy(g["recon"])).abs().max()) <= 1e-6
    with pytest.raises(RuntimeError):
        m(x)                                  # no CPU path
    with pytest.raises(RuntimeError):
        Decoder()(x)                          # container only


def test_clipunet_prompt_tree_and_oracle_against_reference_golden():
    """ClipUnetPrompt drop-in (reference models/prompt_segmentation.py:16-95): same keys/shapes/seeded init as the
    reference for everything but the CLIP encoder, and the oracle (PromptEncoder + prompt_fusion restated) reproduces the
    reference's logits on the golden images, heat-maps and embedding."""
    from transformers import CLIPConfig, CLIPModel
    from image_segmentation_b200 import synthetic
    from image_segmentation_b200.models.prompt_segmentation import ClipUnetPrompt, PromptEncoder
    from oracle.make_golden import state_checksum
    g = np.load(os.path.join(GOLD, "clipunet_prompt_small.npz"))
    torch.manual_seed(0)
    m = ClipUnetPrompt(clip_model=CLIPModel(CLIPConfig()).eval()).eval()
    own = {k: v for k, v in m.state_dict().items() if not k.startswith("clip_feature_extractor.")}
    assert list(own.keys()) == [str(k) for k in g["keys"]]
    assert [str(tuple(v.shape)) for v in own.values()] == [str(s) for s in g["shapes"]]
    assert state_checksum(own) == str(g["seed42_state_sha256"])
    sd = synthetic.amplify_bottleneck_path(synthetic.randomize_batchnorm(own, seed=7))
    assert state_checksum(sd) == str(g["final_state_sha256"])
    x = synthetic.to_float_nchw(torch.from_numpy(g["images_u8"]))
    heat = torch.from_numpy(g["heatmaps"])
    assert torch.equal(heat, synthetic.prompt_heatmaps(2, 32, 64, seed=3, sigma=3.0))
    ref, ref_flip = torch.from_numpy(g["logits"]), torch.from_numpy(g["logits_flipped_prompt"])
    assert ref.shape == (2, 1, 32, 64)
    assert float((ref - ref_flip).abs().max()) > 0.1 * float(ref.abs().max())   # the prompt branch matters in this fixture
    for h, want in ((heat, ref), (heat.flip(-1).contiguous(), ref_flip)):
        logits = unet_oracle.forward(sd, x, clip_feats=torch.from_numpy(g["clip_feats"]), prompt=h)
        assert float((logits - want).abs().max()) <= 2e-5 * float(want.abs().max())
    with pytest.raises(RuntimeError):
        m(x, heat)                            # no CPU path
    with pytest.raises(RuntimeError):
        PromptEncoder()(heat)                 # runs only inside the model


def test_dice_restatement_is_self_consistent():
    """Dice is PARITY UNPINNED (smp absent): only internal consistency can be checked."""
    g = torch.Generator().manual_seed(0)
    z = torch.randn((2, 3, 16, 16), generator=g)
    t = torch.randint(0, 3, (2, 16, 16), generator=g)
    d = mo.dice_smp(z, t)
    assert 0.0 < d < 1.0
    t2 = torch.zeros_like(t)  # classes 1,2 absent -> masked out of the loss
    s = mo.dice_sums(z, t2).numpy()
    assert s[1, 2] == 0 and s[2, 2] == 0 and mo.dice_from_sums(s) > 2.0 / 3.0


# ------------------------------------------------------------------ live reference (build container)
@pytest.mark.reference
def test_live_reference_unet_forward():
    ns = refload.load()
    for cls, hw in ((ns.UNet, (32, 64)), (ns.LargeUNet, (32, 32))):
        torch.manual_seed(3)
        m = cls().eval()
        x = torch.rand(2, 3, *hw)
        with torch.no_grad():
            ref = m(x)
        assert torch.equal(unet_oracle.forward(m.state_dict(), x), ref)


@pytest.mark.reference
def test_live_reference_metrics():
    ns = refload.load()
    g = torch.Generator().manual_seed(1)
    for b in (1, 8, 10, 16):
        z = torch.randn((b, 3, 32, 32), generator=g)
        t = torch.randint(0, 3, (b, 32, 32), generator=g)
        if b == 1:
            t[t == 2] = 0  # a class absent from the targets
        cm = mo.confusion_matrix(unet_oracle.hard_prediction(z).numpy(), t.numpy())
        assert float(ns.losses.IoU()(z, t)) == float(mo.iou_from_cm(cm))
        assert float(ns.losses.PixelAccuracy()(z, t)) == float(mo.pixel_accuracy_from_cm(cm))


@pytest.mark.reference
def test_live_reference_wrapper_levels():
    """All 79 TestWrapper levels (model_wrappers.py:743-764) of the float-space modules, bit-exact."""
    ns = refload.load()
    x = torch.rand((3, 3, 20, 28), generator=torch.Generator().manual_seed(2))
    xn = x.numpy()
    for std in [1e-6, 2, 4, 6, 8, 10, 12, 14, 16, 18]:
        torch.manual_seed(5); ref = ns.blocks.GaussianPixelNoise(std)(x.clone())
        torch.manual_seed(5); aux = torch.randn_like(x).numpy()
        assert np.array_equal(po.f_gaussian_noise(xn, aux, std), ref.numpy())
    for f in [1.0, 1.01, 1.02, 1.03, 1.04, 1.05, 1.1, 1.15, 1.20, 1.25, 0.95, 0.90, 0.85, 0.80, 0.60, 0.40, 0.30, 0.20, 0.10]:
        assert np.array_equal(po.f_contrast(xn, f), ns.blocks.ContrastChange(f)(x.clone()).numpy())
    for o in [0, 5, 10, 15, 20, 25, 30, 35, 40, 45, -5, -10, -15, -20, -25, -30, -35, -40, -45]:
        assert np.array_equal(po.f_brightness(xn, o), ns.blocks.BrightnessChange(o)(x.clone()).numpy())
    for s in [0, 5, 10, 15, 20, 25, 30, 35, 40, 45]:
        random.seed(9); ref = ns.blocks.Occlusion(s)(x.clone())
        random.seed(9); corners = po.replay_f_occlusion_corners(3, s, 20, 28)
        assert np.array_equal(po.f_occlusion(xn, s, corners), ref.numpy())
    for a in [0.00, 0.02, 0.04, 0.06, 0.08, 0.10, 0.12, 0.14, 0.16]:
        torch.manual_seed(6); ref = ns.blocks.SaltAndPepper(a)(x.clone())
        torch.manual_seed(6); r = torch.rand((3, 1, 20, 28)).numpy()
        assert np.array_equal(po.f_salt_pepper(xn, r, a), ref.numpy())
