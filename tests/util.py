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


def train_like(sd, images_f32, masks, steps=150, lr=2e-3, batch=8, crop=128, seed=0, device=None, clip_feats=None,
               prompts=None, binary=False, augment=False):
    """A few Adam steps of cross-entropy on the oracle network (BatchNorm kept in eval mode = fixed
    affine) so the logits get real margins.  Checker-side only; runs on the GPU box's torch (IEEE fp32, conftest.py).
    clip_feats [n,512] / prompts [n,1,H,W]: the ClipUnet / ClipUnetPrompt variants of the oracle network;
    binary: 1-class head trained with BCE against (mask > 0).
    augment: photometric jitter per step (contrast x[0.1,1.3], brightness +-0.2, pixel noise up to 18/255) — a network
    meant for a robustness sweep keeps margins under the sweep's perturbations; without it the extreme levels
    (contrast 0.1: a nearly black image) sit on thin margins where 0.07 % of flipped pixels move a 16-image IoU by
    more than 0.1 point."""
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    # run-to-run repeatable fixtures: deterministic cuDNN kernels, no autotuning (the gates downstream sit on real margins,
    # but a fixture that changes between runs makes a rare thin level a flaky test)
    det = (torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark)
    torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = True, False
    try:
        return _train_like(sd, images_f32, masks, steps, lr, batch, crop, seed, device, clip_feats, prompts, binary, augment)
    finally:
        torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = det


def _train_like(sd, images_f32, masks, steps, lr, batch, crop, seed, device, clip_feats, prompts, binary, augment):
    g = torch.Generator().manual_seed(seed)
    params = {k: v.clone().to(device).requires_grad_(v.is_floating_point() and "running" not in k)
              for k, v in sd.items() if v.is_floating_point()}
    opt = torch.optim.Adam([p for p in params.values() if p.requires_grad], lr=lr)
    n, _, h, w = images_f32.shape
    images_f32, masks = images_f32.to(device), masks.to(device).long()
    if clip_feats is not None:
        clip_feats = clip_feats.to(device)
    if prompts is not None:
        prompts = prompts.to(device)
    for _ in range(steps):
        idx = torch.randint(0, n, (batch,), generator=g)
        y0 = int(torch.randint(0, h - crop + 1, (1,), generator=g))
        x0 = int(torch.randint(0, w - crop + 1, (1,), generator=g))
        xb = images_f32[idx][:, :, y0:y0 + crop, x0:x0 + crop]
        tb = masks[idx][:, y0:y0 + crop, x0:x0 + crop]
        if augment:
            u = torch.rand((4,), generator=g)
            # a third of the steps sit at the dark end of the sweep (contrast 0.05..0.25): that is where margins are thinnest
            factor = float(0.05 + 0.2 * u[0]) if float(u[3]) < 0.33 else float(0.1 + 1.2 * u[0])
            xb = xb * factor + float(0.4 * u[1] - 0.2) * min(1.0, 2.0 * factor)
            xb = (xb + torch.randn(xb.shape, generator=g).to(device) * float(u[2]) * (18.0 / 255.0)).clamp(0, 1)
        kw = {}
        if clip_feats is not None:
            kw["clip_feats"] = clip_feats[idx.to(device)]
        if prompts is not None:
            kw["prompt"] = prompts[idx.to(device)][:, :, y0:y0 + crop, x0:x0 + crop]
        with torch.enable_grad():
            z = unet_oracle.forward_impl(params, xb, **kw)
            if binary:
                loss = torch.nn.functional.binary_cross_entropy_with_logits(z[:, 0], (tb > 0).float())
            else:
                loss = torch.nn.functional.cross_entropy(z, tb)
            opt.zero_grad()
            loss.backward()
            opt.step()
    out = {k: v.detach().cpu() for k, v in params.items()}
    for k, v in sd.items():
        if k not in out:
            out[k] = v.clone()
    return out


_announced = False


def oracle_logits(sd, x, **kw):
    """The oracle network evaluated ON THE CPU in fp32 (never cuDNN/TF32): the reference side of every logits / argmax
    gate.  Inputs may live anywhere; the result is a CPU tensor."""
    global _announced
    if not _announced:
        print(f"[oracle] unet_oracle.forward on the CPU, fp32, {torch.get_num_threads()} threads")
        _announced = True
    sd_cpu = {k: v.detach().cpu() for k, v in sd.items()}
    kw_cpu = {k: (v.detach().cpu() if torch.is_tensor(v) else v) for k, v in kw.items()}
    return unet_oracle.forward(sd_cpu, x.detach().cpu(), **kw_cpu)


def oracle_logits_gpu_ieee(sd, x, dev, **kw):
    """The oracle on the GPU with TF32 switched off (tests/conftest.py): IEEE fp32 through cuDNN.  Used only where the
    CPU would take minutes (the 159-level sweep); test_gpu_sweep_parity.py checks it against oracle_logits."""
    assert not torch.backends.cudnn.allow_tf32 and not torch.backends.cuda.matmul.allow_tf32
    kw_dev = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in kw.items()}
    return unet_oracle.forward({k: v.to(dev) for k, v in sd.items()}, x.to(dev), **kw_dev).cpu()


def argmax_agreement(a: torch.Tensor, b: torch.Tensor) -> float:
    """Share of pixels where the reference rule argmax(softmax(z)) agrees; for a 1-class head, where sign(z) agrees."""
    if a.shape[1] == 1:
        return float(((a > 0) == (b > 0)).float().mean())
    return float((unet_oracle.hard_prediction(a) == unet_oracle.hard_prediction(b)).float().mean())
