"""fp32-mode gradient discrepancy hunt: (n_classes, bilinear) grid x {our loss, oracle dlogits injected}."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import semantic_segmentation_b200 as S
from oracle import genseg_oracle as O
from semantic_segmentation_b200.unet import UNet
from semantic_segmentation_b200.util.dice_score import seg_loss

def rel(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))

S.set_precision("fp32")
for ncls in (1, 2):
    for bil in (True, False):
        st = O.seeded_state(O.unet_state_spec(1, ncls, bil), 3)
        x, y = O.seeded_inputs(4, 2, 1, 32, 32, ncls)
        P = {k: v.clone().requires_grad_(True) for k, v in st.items() if v.dtype.is_floating_point and "running" not in k}
        wk = dict(st); wk.update(P)
        lg = O.unet_forward(wk, x, bil, True); lg.retain_grad()
        loss, _, _ = O.seg_loss(lg, y); loss.backward()
        net = UNet(1, ncls, bil); net.load_state_dict(st); net = net.cuda().train()
        for inject in (False, True):
            net.zero_grad(set_to_none=True)
            out = net(x.cuda())
            if inject:
                out.backward(lg.grad.cuda())
            else:
                seg_loss(out, y.cuda()).backward()
            errs = {k: rel(p.grad, P[k].grad) for k, p in net.named_parameters()}
            worst = max(errs, key=errs.get)
            last = "up4.conv.double_conv.4.bias"
            print("ncls=%d bilinear=%d inject_ref_dlogits=%d: logits %.2e worst %.2e (%s) last-bn-bias %.2e outc.w %.2e"
                  % (ncls, bil, inject, rel(out, lg), errs[worst], worst, errs[last], errs["outc.conv.weight"]), flush=True)
