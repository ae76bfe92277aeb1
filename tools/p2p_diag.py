"""fp32 validation mode of the Pix2Pix generator: per-layer forward error (normalised pre-activations of every down / up level)
of the GPU path and of the CPU fp32 oracle, both against the fp64 oracle.  Usage: python tools/p2p_diag.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import semantic_segmentation_b200 as S
from oracle import genseg_oracle as O
from semantic_segmentation_b200.models_pix2pix import networks


def rel(a, b):
    a, b = a.detach().double().cpu().flatten(), b.detach().double().cpu().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


gold = np.load(os.path.join(ROOT, "tests", "golden", "pix2pix_drift.npz"))
st = O.seeded_state(O.generator_state_spec(1, 1, 64, 8), 11, conv_std=0.02)
arch = torch.from_numpy(gold["arch"])
g = torch.Generator().manual_seed(14)
mask = torch.randint(0, 2, (2, 1, 256, 256), generator=g).float()
t64, t32 = {}, {}
f64 = O.generator_forward({k: (v.double() if v.dtype.is_floating_point else v) for k, v in st.items()}, mask.double(), arch.double(), True, taps=t64)
f32 = O.generator_forward(st, mask, arch, True, taps=t32)
S.set_precision("fp32")
G = networks.define_G(1, 1, 64, "unet_256", "batch", False, "normal", 0.02, [])
G.load_state_dict(st, strict=True)
G = G.cuda().train()
P = {k: v.detach() for k, v in G.named_parameters()}
P["__upconv_arch__"] = arch.cuda()
B = {k: v.clone() for k, v in G.named_buffers()}
with torch.no_grad():
    out, tape = G._program.forward(mask.cuda(), P, B, True, True)
print("fake: gpu-fp32 vs fp64 %.3e   cpu-fp32 vs fp64 %.3e" % (rel(out, f64), rel(f32, f64)))
for name in ["down%d" % d for d in range(1, 8)] + ["up%d" % d for d in range(7, 0, -1)]:
    d = int(name[-1])
    sv = tape[("down%d" % (d + 1)) if name.startswith("down") else ("up%d" % d)]
    y = sv.y.float()
    if sv.scale is not None:
        c = sv.center if sv.center is not None else torch.zeros_like(sv.scale)
        z = (y - c) * sv.scale + sv.shift
    else:
        z = y
    z = z.permute(0, 3, 1, 2)
    raw_ref = None
    print("%-6s z: gpu-fp32 vs fp64 %.3e   cpu-fp32 vs fp64 %.3e   (shape %s)" % (name, rel(z, t64[name]), rel(t32[name], t64[name]), tuple(z.shape)))
