"""Shared helpers for the parity tests (test infrastructure; may use torch/cuDNN as a checker)."""
import numpy as np
import torch

from oracle import unet_oracle


def bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).float()


def nchw_to_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    return x.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def nhwc_to_nchw_f32(x: torch.Tensor) -> torch.Tensor:
    return x.float().permute(0, 3, 1, 2).contiguous()


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """max |a-b| / max |b| (scale-relative, the north_star's 'logits within 1e-2 relative')."""
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def centre_head(sd, x):
    """out.bias := -mean logit per class on x, so all classes appear (SURVEY 8d recipe 2)."""
    sd = dict(sd)
    logits = unet_oracle.forward(sd, x)
    sd["out.bias"] = sd["out.bias"] - logits.mean(dim=(0, 2, 3))
    return sd


def train_like(sd, images_f32, masks, steps=150, lr=2e-3, batch=8, crop=128, seed=0, device=None):
    """A few Adam steps of cross-entropy on the oracle network (BatchNorm kept in eval mode = fixed
    affine) so the logits get real margins.  Checker-side only; runs on the GPU box's torch."""
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    g = torch.Generator().manual_seed(seed)
    params = {k: v.clone().to(device).requires_grad_(v.is_floating_point() and "running" not in k)
              for k, v in sd.items() if v.is_floating_point()}
    opt = torch.optim.Adam([p for p in params.values() if p.requires_grad], lr=lr)
    n, _, h, w = images_f32.shape
    images_f32, masks = images_f32.to(device), masks.to(device).long()
    for _ in range(steps):
        idx = torch.randint(0, n, (batch,), generator=g)
        y0 = int(torch.randint(0, h - crop + 1, (1,), generator=g))
        x0 = int(torch.randint(0, w - crop + 1, (1,), generator=g))
        xb = images_f32[idx][:, :, y0:y0 + crop, x0:x0 + crop]
        tb = masks[idx][:, y0:y0 + crop, x0:x0 + crop]
        with torch.enable_grad():
            loss = torch.nn.functional.cross_entropy(unet_oracle.forward_impl(params, xb), tb)
            opt.zero_grad()
            loss.backward()
            opt.step()
    out = {k: v.detach().cpu() for k, v in params.items()}
    for k, v in sd.items():
        if k not in out:
            out[k] = v.clone()
    return out
