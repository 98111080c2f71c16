"""TEST INFRASTRUCTURE ONLY (oracle): CPU fp32 restatement of the reference's CLIP image-embedding path,
models/processing_blocks.py:156-233.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.

  * `preprocess`      — CustomClipPreprocessor.forward (:163-170): per image `transforms.Resize((224,224))` then
                        `transforms.Normalize(mean, std)`.  On a tensor torchvision's Resize is
                        F.interpolate(mode="bilinear", antialias=True, align_corners=False); restated with that call so the
                        oracle does not need torchvision.  Pinned: tests/test_oracle_pinned.py compares it bit for bit with
                        torchvision's own transforms (installed in this image, 0.26) and, where /root/reference is present,
                        with the reference's class itself.
  * `image_features`  — ClipFeatureExtractor.forward (:219-233): `clip_model.get_image_features(pixel_values=...)` of
                        HuggingFace transformers' CLIPModel (third-party, transformers 5.5 in this image; the reference pins
                        4.x in requirements.txt — same module tree and arithmetic for the vision tower) in fp32 on the CPU.
  * `vit_restated`    — the same vision tower written out with plain torch ops from the state_dict (embeddings, pre_layrnorm,
                        pre-norm encoder layers with QuickGELU, post_layernorm, visual_projection), pinned against
                        `image_features`; it documents the arithmetic the library implements.
  * `seeded_clip`     — a seeded random-init CLIPModel (the openai/clip-vit-base-patch32 checkpoint needs the network) whose
                        LayerNorm affines and biases are randomised so every parameter of the tower matters.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

CLIP_MEAN = (0.48145466, 0.4578275, 0.40821073)   # processing_blocks.py:204
CLIP_STD = (0.26862954, 0.26130258, 0.27577711)   # processing_blocks.py:205


def preprocess(x: torch.Tensor, size: int = 224) -> torch.Tensor:
    """[n,3,H,W] fp32 -> [n,3,size,size]; processing_blocks.py:163-170 (loop over images, stack)."""
    out = []
    mean = torch.tensor(CLIP_MEAN, dtype=x.dtype).view(3, 1, 1)
    std = torch.tensor(CLIP_STD, dtype=x.dtype).view(3, 1, 1)
    for img in x:
        r = F.interpolate(img[None], size=(size, size), mode="bilinear", antialias=True, align_corners=False)[0]
        out.append((r - mean) / std)
    return torch.stack(out)


def image_features(clip_model, x: torch.Tensor) -> torch.Tensor:
    """processing_blocks.py:219-233 on the CPU in fp32."""
    clip_model = clip_model.float().cpu().eval()
    with torch.no_grad():
        f = clip_model.get_image_features(pixel_values=preprocess(x.float().cpu(), clip_model.config.vision_config.image_size))
    if not torch.is_tensor(f):
        f = f.pooler_output
    return f.float()


def vit_restated(sd: dict, pixel_values: torch.Tensor, heads: int, eps: float = 1e-5) -> torch.Tensor:
    """CLIPModel.get_image_features from the state_dict alone (fp32)."""
    g = lambda k: sd[k].float()
    w = g("vision_model.embeddings.patch_embedding.weight")
    d, p = w.shape[0], w.shape[2]
    x = F.conv2d(pixel_values, w, stride=p)                                # [n, D, G, G], no bias
    x = x.flatten(2).transpose(1, 2)                                       # [n, G*G, D]
    cls = g("vision_model.embeddings.class_embedding").expand(x.shape[0], 1, d)
    x = torch.cat([cls, x], dim=1) + g("vision_model.embeddings.position_embedding.weight")[None]
    x = F.layer_norm(x, (d,), g("vision_model.pre_layrnorm.weight"), g("vision_model.pre_layrnorm.bias"), eps)
    layer = 0
    while f"vision_model.encoder.layers.{layer}.self_attn.q_proj.weight" in sd:
        pre = f"vision_model.encoder.layers.{layer}."
        h = F.layer_norm(x, (d,), g(pre + "layer_norm1.weight"), g(pre + "layer_norm1.bias"), eps)
        q = F.linear(h, g(pre + "self_attn.q_proj.weight"), g(pre + "self_attn.q_proj.bias"))
        k = F.linear(h, g(pre + "self_attn.k_proj.weight"), g(pre + "self_attn.k_proj.bias"))
        v = F.linear(h, g(pre + "self_attn.v_proj.weight"), g(pre + "self_attn.v_proj.bias"))
        n, t, _ = q.shape
        split = lambda z: z.view(n, t, heads, d // heads).transpose(1, 2)   # [n, heads, t, 64]
        a = torch.softmax(split(q) @ split(k).transpose(-1, -2) / math.sqrt(d // heads), dim=-1) @ split(v)
        a = a.transpose(1, 2).reshape(n, t, d)
        x = x + F.linear(a, g(pre + "self_attn.out_proj.weight"), g(pre + "self_attn.out_proj.bias"))
        h = F.layer_norm(x, (d,), g(pre + "layer_norm2.weight"), g(pre + "layer_norm2.bias"), eps)
        h = F.linear(h, g(pre + "mlp.fc1.weight"), g(pre + "mlp.fc1.bias"))
        h = h * torch.sigmoid(1.702 * h)                                    # QuickGELU
        x = x + F.linear(h, g(pre + "mlp.fc2.weight"), g(pre + "mlp.fc2.bias"))
        layer += 1
    pooled = F.layer_norm(x[:, 0], (d,), g("vision_model.post_layernorm.weight"), g("vision_model.post_layernorm.bias"), eps)
    return F.linear(pooled, g("visual_projection.weight"))


def seeded_clip(seed: int = 0, small: bool = False):
    """Random-init CLIPModel (ViT-B/32 by default; `small`: hidden 256, 2 layers, MLP 512 — a second shape for the
    library's dispatch) with LayerNorm affines and all biases of the vision tower randomised."""
    from transformers import CLIPConfig, CLIPModel
    torch.manual_seed(seed)
    if small:
        cfg = CLIPConfig(vision_config=dict(hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                                            num_attention_heads=4, patch_size=32, image_size=224), projection_dim=128)
    else:
        cfg = CLIPConfig()
    m = CLIPModel(cfg).eval()
    gen = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if not (name.startswith("vision_model.") or name.startswith("visual_projection")):
                continue
            if "norm" in name and name.endswith(".weight"):
                p.copy_(0.5 + torch.rand(p.shape, generator=gen))
            elif name.endswith(".bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=gen))
            elif "class_embedding" in name or "position_embedding" in name:
                p.copy_(0.5 * torch.randn(p.shape, generator=gen))
            elif name.endswith(".weight") and p.dim() >= 2:
                # default init leaves the branches tiny next to the residual stream; widen them so attention and MLP matter
                fan_in = p[0].numel()
                p.copy_(torch.randn(p.shape, generator=gen) * (1.0 / math.sqrt(fan_in)))
    return m
