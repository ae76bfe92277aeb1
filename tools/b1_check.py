"""One-GPU diagnostic: U-Net train step at small per-GPU batches (1, 2, 3, 8 images of 64x64) against the CPU
oracle, per-parameter gradient errors.  python tools/b1_check.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import semantic_segmentation_b200 as S
from oracle import genseg_oracle as O
from semantic_segmentation_b200.unet import UNet
from semantic_segmentation_b200.util.dice_score import seg_loss


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


for prec in ("fp32", "bf16"):
    S.set_precision(prec)
    for n, hw in ((1, 64), (2, 64), (3, 64), (8, 64), (1, 256)):
        st = O.seeded_state(O.unet_state_spec(1, 2, True), 1)
        x, y = O.seeded_inputs(2, n, 1, hw, hw, 2)
        ref = {k: v.clone().requires_grad_(v.dtype.is_floating_point and "running" not in k) for k, v in st.items()}
        taps = {}
        lg = O.unet_forward(ref, x, True, True, False, taps)
        tot, _, _ = O.seg_loss(lg, y)
        tot.backward()
        net = UNet(1, 2, True); net.load_state_dict(st); net = net.cuda().train()
        out = net(x.cuda()); loss = seg_loss(out, y.cuda()); loss.backward()
        errs = sorted(((rel(p.grad, ref[k].grad), k) for k, p in net.named_parameters()), reverse=True)
        print("%s n=%d hw=%d: logits %.2e loss %.7f ref %.7f | worst grads: %s" % (
            prec, n, hw, rel(out, lg), loss.item(), tot.item(),
            ", ".join("%s %.1e" % (k, e) for e, k in errs[:4])), flush=True)
