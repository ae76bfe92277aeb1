"""Per-layer forward / gradient error of the U-Net path vs the CPU oracle, for several execution modes.
Usage: python tools/unet_diag.py nch ncls bilinear n h w [modes...]   modes: fp32 bf16_direct bf16_tc_nofuse bf16_tc"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import semantic_segmentation_b200 as S  # noqa: E402
from oracle import genseg_oracle as O  # noqa: E402
from semantic_segmentation_b200.unet import UNet  # noqa: E402
from semantic_segmentation_b200.util.dice_score import seg_loss  # noqa: E402


def rel(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def main():
    nch, ncls, bil, n, h, w = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3] == "1", int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
    modes = sys.argv[7:] or ["fp32", "bf16_direct", "bf16_tc_nofuse", "bf16_tc"]
    st = O.seeded_state(O.unet_state_spec(nch, ncls, bil), 1)
    x, y = O.seeded_inputs(2, n, nch, h, w, ncls)
    params = {k: v.clone().requires_grad_(True) for k, v in st.items() if v.dtype.is_floating_point and "running" not in k}
    work = dict(st)
    work.update(params)
    taps = {}
    ref_logits = O.unet_forward(work, x, bil, training=True, taps=taps)
    ref_loss, _, _ = O.seg_loss(ref_logits, y)
    ref_loss.backward()
    for mode in modes:
        S.set_precision("fp32" if mode == "fp32" else "bf16")
        S.set_tensor_core_mode(0 if mode in ("fp32", "bf16_direct") else 2, fused_stats=(mode == "bf16_tc"))
        net = UNet(nch, ncls, bil)
        net.load_state_dict(st)
        net = net.cuda().train()
        # forward taps through the program (raw conv outputs)
        pd = {k: v.detach() for k, v in net.named_parameters()}
        bd = {k: v.clone() for k, v in net.named_buffers()}
        with torch.no_grad():
            _, tape = net._program.forward(x.cuda(), pd, bd, True, True)
        print("== mode", mode)
        for name in ["inc"] + [f"down{i}" for i in range(1, 5)] + [f"up{i}" for i in range(1, 5)]:
            s0, s1 = tape[name]
            pref = {"inc": "inc"}.get(name, (name + ".maxpool_conv.1") if name.startswith("down") else name + ".conv")
            for sv, idx in ((s0, 0), (s1, 3)):
                ref = taps[f"{pref}.double_conv.{idx}"]
                print("  fwd %-34s %.3e" % (f"{pref}.double_conv.{idx}", rel(sv.y.float().permute(0, 3, 1, 2), ref)))
        logits = net(x.cuda())
        loss = seg_loss(logits, y.cuda())
        loss.backward()
        print("  logits %.3e  loss %.6f ref %.6f" % (rel(logits, ref_logits), loss.item(), ref_loss.item()))
        for k, p in reversed(list(net.named_parameters())):
            print("  grad %-44s %.3e" % (k, rel(p.grad, params[k].grad)))
    S.set_precision("bf16")
    S.set_tensor_core_mode(2, True)


main()
