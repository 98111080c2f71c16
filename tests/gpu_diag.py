"""GPU diagnostics (run on the B200 box; lives under tests/ because its parity modes use the oracle): per-layer and end-to-end parity numbers, printed rather than
asserted, so one gpurun call tells where a kernel is wrong.  Checker side uses torch on the GPU."""
import sys
import time
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from image_segmentation_b200 import synthetic
from image_segmentation_b200.engine import Engine, Perturbation
from image_segmentation_b200.models.UNet import UNet, LargeUNet
from oracle import unet_oracle, metrics_oracle
from tests import util

dev = torch.device("cuda:0")


def make_model(large=False, seed=42):
    torch.manual_seed(seed)
    m = (LargeUNet if large else UNet)().eval()
    sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7)
    m.load_state_dict(sd)
    return m, sd


def folded(sd, conv_key):
    base, idx = conv_key.rsplit(".", 1)
    bn = f"{base}.{int(idx) + 1}"
    s = (sd[bn + ".weight"].double() / torch.sqrt(sd[bn + ".running_var"].double() + 1e-5))
    w = (sd[conv_key + ".weight"].double() * s[:, None, None, None]).float()
    b = ((sd[conv_key + ".bias"].double() - sd[bn + ".running_mean"].double()) * s + sd[bn + ".bias"].double()).float()
    return w, b


def stage_layers(large=False, only=None):
    m, sd = make_model(large)
    eng = m.to(dev).engine()
    g = torch.Generator().manual_seed(3)
    for key, cout in m.layer_couts().items():
        if only and only not in key:
            continue
        for (n, h, w) in [(2, 24, 40), (3, 32, 32)]:
            if key.endswith(".up"):
                cin = sd[key + ".weight"].shape[0]
                x = torch.randn((n, cin, h, w), generator=g)
                xb = util.bf16_round(x)
                ref = F.conv_transpose2d(xb.to(dev), util.bf16_round(sd[key + ".weight"]).to(dev), sd[key + ".bias"].to(dev), stride=2)
                out = eng.debug_convt2x2(key, util.nchw_to_nhwc_bf16(x).to(dev))
                torch.cuda.synchronize()
                got = util.nhwc_to_nchw_f32(out)
                print(f"convT {key:28s} n{n} {h}x{w} cin{cin} cout{cout}: rel_err {util.rel_err(got, ref):.3e}", flush=True)
            else:
                wf, bf = folded(sd, key)
                cin = wf.shape[1]
                x = torch.randn((n, cin, h, w), generator=g)
                xb = util.bf16_round(x)
                ref = F.relu(F.conv2d(xb.to(dev), util.bf16_round(wf).to(dev), bf.to(dev), padding=1))
                for pool in (False, True):
                    r = F.max_pool2d(ref, 2, 2) if pool else ref
                    out = eng.debug_conv3x3(key, util.nchw_to_nhwc_bf16(x).to(dev), pool=pool)
                    torch.cuda.synchronize()
                    got = util.nhwc_to_nchw_f32(out)
                    print(f"conv  {key:28s} n{n} {h}x{w} cin{cin} cout{cout} pool{int(pool)}: rel_err {util.rel_err(got, r):.3e}", flush=True)


def stage_forward(large=False, hw=(64, 128), n=3):
    m, sd = make_model(large)
    imgs, masks = synthetic.make_dataset(n, hw[0], hw[1], seed=5)
    x = synthetic.to_float_nchw(imgs)
    sd = util.centre_head(sd, x)
    m.load_state_dict(sd)
    m = m.to(dev)
    ref, acts = unet_oracle.forward({k: v.to(dev) for k, v in sd.items()}, x.to(dev), return_intermediates=True)
    got = m(x.to(dev))
    torch.cuda.synchronize()
    print(f"forward {'large' if large else 'unet'} {hw}: logits rel_err {util.rel_err(got, ref):.3e}", flush=True)
    pa = unet_oracle.hard_prediction(ref)
    pb = unet_oracle.hard_prediction(got)
    print(f"  argmax agreement {(pa == pb).float().mean().item():.5f}; class hist {torch.bincount(pa.flatten(), minlength=3).tolist()}", flush=True)
    eng = m.engine()
    res = eng.eval_batches(imgs.to(dev), masks.to(dev), None, ref_batch=2, want_dice=True, want_pred=True, want_logits=True)
    torch.cuda.synchronize()
    print(f"  eval logits vs forward logits max diff {(res['logits'] - got).abs().max().item():.3e}")
    pred = res["pred"].cpu().numpy()
    print(f"  pred == argmax(softmax(own logits)): {(torch.from_numpy(pred).to(dev) == pb).float().mean().item():.6f}")
    for bi in range(res["cm"].shape[0]):
        cm_ref = metrics_oracle.confusion_matrix(pred[2 * bi:2 * bi + 2], masks[2 * bi:2 * bi + 2].numpy())
        print(f"  batch {bi}: cm equal {bool((res['cm'][bi].cpu().numpy() == cm_ref).all())}  sum {int(res['cm'][bi].sum())}")
        ds = metrics_oracle.dice_sums(res["logits"][2 * bi:2 * bi + 2].cpu(), masks[2 * bi:2 * bi + 2].long())
        print(f"    dice sums max rel diff {float(((res['dice'][bi].cpu() - ds).abs() / ds.abs().clamp_min(1)).max()):.3e}")


def stage_time(mb=32, n=64, steps=3):
    m, sd = make_model()
    m = m.to(dev)
    eng = m.engine(max_microbatch=mb)
    imgs, masks = synthetic.make_dataset(n, 256, 256, seed=5)
    imgs, masks = imgs.to(dev), masks.to(dev)
    for i in range(steps + 1):
        torch.cuda.synchronize()
        t0 = time.time()
        eng.eval_batches(imgs, masks, Perturbation.u8_contrast(1.1), ref_batch=8)
        torch.cuda.synchronize()
        dt = time.time() - t0
        print(f"eval {n} imgs mb{mb}: {dt * 1e3:.2f} ms -> {n / dt:.1f} img/s", flush=True)


if __name__ == "__main__":
    stage = sys.argv[1]
    if stage == "layers":
        stage_layers(only=sys.argv[2] if len(sys.argv) > 2 else None)
    elif stage == "layers_large":
        stage_layers(large=True, only=sys.argv[2] if len(sys.argv) > 2 else None)
    elif stage == "forward":
        stage_forward()
        stage_forward(hw=(256, 256), n=2)
    elif stage == "forward_large":
        stage_forward(large=True, hw=(128, 128), n=2)
    elif stage == "time":
        stage_time(mb=int(sys.argv[2]) if len(sys.argv) > 2 else 32, n=int(sys.argv[3]) if len(sys.argv) > 3 else 64)
