"""The CLIP image-embedding path in the library (`ub_clip_image_features`, UB_ARCH_CLIP_VIT) against the CPU fp32 oracle
(oracle/clip_oracle.py = torchvision-equivalent Resize + Normalize, then HuggingFace CLIPModel.get_image_features):
models/processing_blocks.py:156-233.

Bars: the preprocessing (fp32 both sides, different summation order) within 1e-5 absolute on values of magnitude <= 2.7;
the embedding (bf16 tensor-core GEMMs, fp32 everything else, 12 layers) within 1e-2 relative (north_star's bf16 bar) with
cosine similarity per image >= 0.9999.  Weights: seeded random-init towers whose LayerNorm affines, biases and branch
weights are widened so attention, MLP and every bias matter (the real checkpoint needs the network).
"""
import pytest
import torch

from oracle import clip_oracle as co
from tests import util

pytestmark = pytest.mark.gpu


def _images(n, h, w, seed):
    g = torch.Generator().manual_seed(seed)
    low = torch.rand((n, 3, 8, 8), generator=g)
    x = torch.nn.functional.interpolate(low, size=(h, w), mode="bicubic", align_corners=False)
    x = (x + 0.08 * torch.randn((n, 3, h, w), generator=g)).clamp(0, 1)
    return (x * 255).round() / 255


@pytest.fixture(scope="module")
def towers(dev):
    from image_segmentation_b200.engine import Engine
    out = {}
    for name, small in (("b32", False), ("small", True)):
        m = co.seeded_clip(seed=3, small=small)
        eng = Engine(dev, "clip_vit", 3, 1, 16)
        eng.load_state_dict({k: v for k, v in m.state_dict().items()
                             if k.startswith("vision_model.") or k == "visual_projection.weight"})
        out[name] = (m, eng)
    yield out
    for _, eng in out.values():
        eng.close()


@pytest.mark.parametrize("hw", [(256, 256), (224, 224), (200, 300), (512, 384), (100, 120), (640, 1000)])
def test_clip_preprocess_matches_resize_normalize(towers, dev, hw):
    """CustomClipPreprocessor (processing_blocks.py:156-170): antialiased bilinear Resize to 224 + Normalize; shrinking,
    identity, mixed and enlarging sizes."""
    _, eng = towers["b32"]
    x = _images(2, hw[0], hw[1], seed=hw[0] + hw[1])
    want = co.preprocess(x)
    got = eng.clip_preprocess(x.to(dev)).cpu()
    err = (got - want).abs().max().item()
    print(f"preprocess {hw} -> 224x224: max abs err {err:.2e}")
    assert got.shape == want.shape and err < 1e-5
    if hw == (224, 224):   # no resampling at the model's own size: only the normalisation
        mean = torch.tensor(co.CLIP_MEAN).view(1, 3, 1, 1)
        std = torch.tensor(co.CLIP_STD).view(1, 3, 1, 1)
        assert (got - (x - mean) / std).abs().max().item() < 1e-6


def test_clip_preprocess_rejects_huge_inputs(towers, dev):
    _, eng = towers["b32"]
    with pytest.raises(RuntimeError, match="taps per axis"):
        eng.clip_preprocess(torch.zeros((1, 3, 1400, 224), device=dev))


@pytest.mark.parametrize("name,n,hw", [("b32", 3, (256, 256)), ("b32", 5, (224, 224)), ("small", 4, (256, 256)),
                                       ("small", 1, (200, 300))])
def test_clip_image_features_vs_fp32_oracle(towers, dev, name, n, hw):
    m, eng = towers[name]
    x = _images(n, hw[0], hw[1], seed=11 + n)
    want = co.image_features(m, x)
    got = eng.clip_image_features(x.to(dev)).cpu()
    rel = util.rel_err(got, want)
    cos = torch.nn.functional.cosine_similarity(got, want, dim=1).min().item()
    # the embeddings must also tell the images apart by far more than the error
    spread = util.rel_err(want.roll(1, 0), want) if n > 1 else float("inf")
    print(f"CLIP tower {name}, {n} x {hw}: embedding rel err {rel:.3e}, min cosine {cos:.6f}, image-to-image spread {spread:.3e}")
    assert got.shape == want.shape
    assert rel < 1e-2 and cos >= 0.9999 and spread > 10 * rel


def test_clip_restated_tower_equals_transformers(towers):
    """The arithmetic the library implements, written out (oracle/clip_oracle.py: vit_restated), is HuggingFace's."""
    for name in ("small", "b32"):
        m, _ = towers[name]
        x = _images(2, 256, 256, seed=5)
        want = co.image_features(m, x)
        cfg = m.config.vision_config
        got = co.vit_restated(m.state_dict(), co.preprocess(x, cfg.image_size), cfg.num_attention_heads, cfg.layer_norm_eps)
        assert util.rel_err(got, want) < 1e-5, name


def test_clip_chunks_and_batch_independence(towers, dev):
    """More images than the context's micro-batch (16): chunks of 16 + a ragged one; every image's embedding equals the
    one it gets alone (the token rows of a GEMM are independent; attention never crosses images)."""
    m, eng = towers["small"]
    x = _images(37, 256, 256, seed=8).to(dev)
    all_ = eng.clip_image_features(x)
    for i in (0, 15, 16, 36):
        alone = eng.clip_image_features(x[i:i + 1].contiguous())
        assert torch.equal(alone[0], all_[i]), i
    assert eng.clip_image_features(x[:0].contiguous()).shape == (0, 128)


def test_clip_feature_extractor_module_backends(dev):
    """The drop-in ClipFeatureExtractor: library backend by default, transformers on request; same embeddings within
    the bf16 bar; unsupported towers are refused by name."""
    from transformers import CLIPConfig, CLIPModel
    from image_segmentation_b200.models.processing_blocks import ClipFeatureExtractor
    m = co.seeded_clip(seed=4, small=True)
    x = _images(3, 256, 256, seed=2).to(dev)
    ours = ClipFeatureExtractor(clip_model=m)
    assert ours.backend == "b200"
    a = ours(x)
    b = ClipFeatureExtractor(clip_model=m, backend="transformers")(x)
    want = co.image_features(m, x.cpu())
    print(f"ClipFeatureExtractor: b200 rel err {util.rel_err(a.cpu(), want):.3e}, transformers (GPU fp32) {util.rel_err(b.cpu(), want):.3e}")
    assert util.rel_err(a.cpu(), want) < 1e-2 and util.rel_err(b.cpu(), want) < 1e-3
    assert ours.vit_engine(dev) is ours.vit_engine(dev)                      # one context, reused
    with pytest.raises(RuntimeError, match="no CPU path"):
        ours(x.cpu())
    gelu = CLIPModel(CLIPConfig(vision_config=dict(hidden_size=256, intermediate_size=512, num_hidden_layers=1,
                                                   num_attention_heads=4, hidden_act="gelu"))).eval()
    with pytest.raises(RuntimeError, match="hidden_act"):
        ClipFeatureExtractor(clip_model=gelu)(x)
