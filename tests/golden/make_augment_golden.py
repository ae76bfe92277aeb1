"""Generates tests/golden/augment_chain.npz by executing the reference's lines
(running_files/train_end2end_jsrt.py:197-200) literally on the CPU with torch + torchvision.
Run from the repo root:  python tests/golden/make_augment_golden.py"""
import os
import numpy as np
import torch
import torchvision.transforms as transforms


def reference_lines(fake_image: torch.Tensor) -> torch.Tensor:
    fake_image = ((fake_image - fake_image.min()) / (fake_image.max() - fake_image.min())).detach()
    fake_image = fake_image.mul(255).add_(0.5).clamp_(0, 255).to(torch.uint8)
    fake_image = transforms.functional.adjust_gamma(transforms.functional.equalize(fake_image), 0.5) / 255.0
    return fake_image


def cases():
    g = torch.Generator().manual_seed(77)
    smooth = torch.tanh(0.7 * torch.randn(3, 1, 40, 56, generator=g))                       # tanh-shaped generator output
    few = (torch.randint(0, 3, (2, 1, 16, 16), generator=g).float() - 1.0) * 0.5               # 3 grey levels
    const_plane = torch.tanh(torch.randn(3, 1, 24, 24, generator=g))
    const_plane[1] = 0.25                                                                      # one constant image: step == 0
    rgb = torch.rand(2, 3, 20, 28, generator=g) * 2 - 1                                        # 3-channel (ISIC) batch
    tiny = torch.tensor([[[[0.0, 1.0], [0.5, 0.25]]]])
    return {"smooth": smooth, "few": few, "const_plane": const_plane, "rgb": rgb, "tiny": tiny}


if __name__ == "__main__":
    out = {}
    for name, x in cases().items():
        out["in:" + name] = x.numpy()
        out["out:" + name] = reference_lines(x.clone()).numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "augment_chain.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items() if k.startswith("out:")})
