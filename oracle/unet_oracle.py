"""ORACLE (test infrastructure, not product): CPU fp32 restatement of the reference U-Net forward.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this.  It restates, as one flat functional pass over a reference-format state_dict:

  models/UNet.py:60-76 (UNet.forward) and :129-148 (LargeUNet.forward)
  models/processing_blocks.py:42-52  ConvBlock  (conv3x3 pad1 + BatchNorm eval + ReLU, twice)
  models/processing_blocks.py:71-77  ConvBlockDownsample (ConvBlock + MaxPool2d(2,2))
  models/CLIP_models.py:114-131 (ClipUnet.forward, given the CLIP embedding) with
  models/processing_blocks.py:310-322 (CrossAttentionFusion)
  models/processing_blocks.py:102-109 ConvBlockUpsampleSkip
        (ConvTranspose2d k2 s2 -> F.interpolate bilinear align_corners=True to skip size ->
         cat([up, skip]) -> ConvBlock)

Pinned: tests/test_oracle_pinned.py compares it with the real reference modules (imported from
/root/reference in the build container) and with tests/golden/unet_small.npz made by
oracle/make_golden.py.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # nn.BatchNorm2d default, processing_blocks.py:44,47


def arch_of(state_dict) -> str:
    """'unet' (3 encoder stages) or 'large' (4), from the keys present."""
    return "large" if any(k.startswith("enc4.") for k in state_dict) else "unet"


def _conv_bn_relu(x, sd, prefix, i_conv, i_bn):
    x = F.conv2d(x, sd[f"{prefix}.{i_conv}.weight"], sd[f"{prefix}.{i_conv}.bias"], padding=1)
    x = F.batch_norm(x, sd[f"{prefix}.{i_bn}.running_mean"], sd[f"{prefix}.{i_bn}.running_var"],
                     sd[f"{prefix}.{i_bn}.weight"], sd[f"{prefix}.{i_bn}.bias"],
                     training=False, eps=BN_EPS)
    return F.relu(x)


def _double_conv(x, sd, prefix):
    x = _conv_bn_relu(x, sd, prefix, 0, 1)
    return _conv_bn_relu(x, sd, prefix, 3, 4)


def _up_skip(x, skip, sd, name):
    x = F.conv_transpose2d(x, sd[f"{name}.up.weight"], sd[f"{name}.up.bias"], stride=2)
    x = F.interpolate(x, size=skip.shape[2:], mode="bilinear", align_corners=True)
    return _double_conv(torch.cat([x, skip], dim=1), sd, f"{name}.conv.conv")


def clip_fusion(sd, clip_feats: torch.Tensor, h: int, w: int) -> torch.Tensor:
    """models/processing_blocks.py:310-322 (CrossAttentionFusion.forward) restated for its actual shapes: every one of
    the H*W queries attends to ONE key/value token per image, so softmax over the key axis is 1 and the output is
    out_proj(v_proj(clip)) at every position, whatever the query.  -> [B, E, h, w].  (tests/test_oracle_pinned.py
    checks this against nn.MultiheadAttention itself.)"""
    e = sd["cross_attention_fusion.cross_attn.out_proj.weight"].shape[0]
    w_v = sd["cross_attention_fusion.cross_attn.in_proj_weight"][2 * e:3 * e]
    b_v = sd["cross_attention_fusion.cross_attn.in_proj_bias"][2 * e:3 * e]
    v = F.linear(clip_feats.float(), w_v, b_v)
    o = F.linear(v, sd["cross_attention_fusion.cross_attn.out_proj.weight"],
                 sd["cross_attention_fusion.cross_attn.out_proj.bias"])
    return o[:, :, None, None].expand(-1, -1, h, w)


def prompt_encoder(sd, heat: torch.Tensor) -> torch.Tensor:
    """models/prompt_segmentation.py:16-29 (PromptEncoder.forward): three pooled double-conv stages, then a ConvBlock."""
    h = heat.float()
    for i in (1, 2, 3):
        h = F.max_pool2d(_double_conv(h, sd, f"prompt_encoder.enc{i}.block.0.conv"), 2, 2)
    return _double_conv(h, sd, "prompt_encoder.conv.conv")


def forward_impl(state_dict, x: torch.Tensor, return_intermediates: bool = False, clip_feats=None, prompt=None):
    """x: fp32 [B,3,H,W] -> logits fp32 [B,C,H,W] (activation = Identity).  Differentiable.
    clip_feats [B,512]: ClipUnet (models/CLIP_models.py:114-131) — the bottleneck handed to dec1 is the fusion.
    prompt [B,1,H,W] (with clip_feats): ClipUnetPrompt (models/prompt_segmentation.py:63-95) — dec1's input is
    prompt_fusion(cat([fusion, prompt_encoder(prompt)]))."""
    sd = {k.replace("_orig_mod.", ""): v.float() for k, v in state_dict.items()
          if v.is_floating_point()}
    n_enc = 4 if arch_of(sd) == "large" else 3
    acts = {}
    stem = F.conv2d(x.float(), sd["input.weight"], sd["input.bias"])
    acts["input"] = stem
    skips = [stem]
    h = stem
    for i in range(1, n_enc + 1):
        h = F.max_pool2d(_double_conv(h, sd, f"enc{i}.block.0.conv"), 2, 2)
        acts[f"enc{i}"] = h
        skips.append(h)
    if clip_feats is not None:
        h = clip_fusion(sd, clip_feats, h.shape[2], h.shape[3])  # CLIP_models.py:122-124: bottleneck(enc3) is unused
        if prompt is not None:  # prompt_segmentation.py:79-86
            cat = torch.cat([h, prompt_encoder(sd, prompt)], dim=1)
            h = F.conv2d(cat, sd["prompt_fusion.weight"], sd["prompt_fusion.bias"])
    else:
        h = _double_conv(h, sd, "bottleneck.conv")
    acts["bottleneck"] = h
    for j in range(1, n_enc + 2):
        h = _up_skip(h, skips[n_enc + 1 - j], sd, f"dec{j}")
        acts[f"dec{j}"] = h
    logits = F.conv2d(h, sd["out.weight"], sd["out.bias"])
    return (logits, acts) if return_intermediates else logits


@torch.no_grad()
def forward(state_dict, x: torch.Tensor, return_intermediates: bool = False, clip_feats=None, prompt=None):
    return forward_impl(state_dict, x, return_intermediates, clip_feats, prompt)


@torch.no_grad()
def autoencoder_forward(state_dict, x: torch.Tensor) -> torch.Tensor:
    """models/autoencoder.py:17-25,36-41,53-58 restated on the flat key names (`encoder.` / `decoder.` prefixes dropped):
    Encoder (stem, enc1-3, bottleneck) -> three ConvBlockUpsample stages (processing_blocks.py:131-133) -> 1x1 -> sigmoid."""
    sd = {k.replace("_orig_mod.", "").replace("encoder.", "", 1).replace("decoder.", "", 1) if k.startswith(("encoder.", "decoder."))
          else k.replace("_orig_mod.", ""): v.float() for k, v in state_dict.items() if v.is_floating_point()}
    h = F.conv2d(x.float(), sd["input.weight"], sd["input.bias"])
    for i in (1, 2, 3):
        h = F.max_pool2d(_double_conv(h, sd, f"enc{i}.block.0.conv"), 2, 2)
    h = _double_conv(h, sd, "bottleneck.conv")
    for j in (1, 2, 3):
        h = F.conv_transpose2d(h, sd[f"dec{j}.up.weight"], sd[f"dec{j}.up.bias"], stride=2)
        h = _double_conv(h, sd, f"dec{j}.conv.conv")
    return torch.sigmoid(F.conv2d(h, sd["out.weight"], sd["out.bias"]))


def hard_prediction(logits: torch.Tensor) -> torch.Tensor:
    """models/losses.py:48-50 — argmax taken AFTER softmax, first index on ties."""
    return torch.argmax(F.softmax(logits, dim=1), dim=1)
