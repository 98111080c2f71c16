"""A/B of the folded stem (default) against the materialised stem (UB_NO_STEM_FOLD=1) and the fp32 oracle."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_segmentation_b200 import synthetic
from image_segmentation_b200.engine import Engine, Perturbation as P
from image_segmentation_b200.models.UNet import UNet
from oracle import unet_oracle
from tests import util
dev = torch.device("cuda:0")
torch.manual_seed(42)
m = UNet().eval(); sd = synthetic.randomize_batchnorm(m.state_dict(), seed=7); m.load_state_dict(sd)
imgs, masks = synthetic.make_dataset(7, 64, 96, seed=3)
x = synthetic.to_float_nchw(imgs).contiguous()
ref = unet_oracle.forward(sd, x)
res = {}
for fold in (True, False):
    if fold: os.environ.pop("UB_NO_STEM_FOLD", None)
    else: os.environ["UB_NO_STEM_FOLD"] = "1"
    eng = Engine(dev, "unet", 3, 3, 3); eng.load_state_dict(sd)
    for name, spec, src in (("none_f32", None, x.to(dev)), ("lut_u8", P.u8_contrast(1.0), imgs.to(dev)), ("blur0_u8", P.u8_blur(0), imgs.to(dev))):
        r = eng.eval_batches(src, masks.to(dev), spec, 4, want_logits=True, want_pred=True)
        torch.cuda.synchronize()
        res[(fold, name)] = r["logits"].cpu()
    eng.close()
for name in ("none_f32", "lut_u8", "blur0_u8"):
    a, b = res[(True, name)], res[(False, name)]
    agree = float((a.argmax(1) == ref.argmax(1)).float().mean()); agree_b = float((b.argmax(1) == ref.argmax(1)).float().mean())
    print(f"{name}: fold vs oracle {util.rel_err(a, ref):.3e} (argmax {agree:.4f}) | materialised vs oracle {util.rel_err(b, ref):.3e} (argmax {agree_b:.4f}) | fold vs materialised {util.rel_err(a, b):.3e}")
