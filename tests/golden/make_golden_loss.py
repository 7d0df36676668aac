"""Golden fixture for the image losses, produced by IMPORTING THE REFERENCE'S utils/loss_utils.py (torch CPU).
Runs only in the build container (needs /root/reference); writes tests/golden/loss_ssim.npz.

    python tests/golden/make_golden_loss.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
from utils.loss_utils import l1_loss, ssim  # noqa: E402  reference

out = {}
for name, (C, H, W), seed in (("a", (3, 40, 56), 0), ("b", (3, 17, 23), 1), ("c", (1, 64, 64), 2)):
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g)
    img = (gt + 0.15 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
    img[:, : H // 3] = gt[:, : H // 3]          # a region with exactly zero error (sign(0) = 0)
    img.requires_grad_(True)
    Ll1 = l1_loss(img, gt)
    s = ssim(img, gt)
    lam = 0.2
    loss = (1.0 - lam) * Ll1 + lam * (1.0 - s)   # train_baseline.py:126-127
    g_l1, = torch.autograd.grad(Ll1, img, retain_graph=True)
    g_ss, = torch.autograd.grad(s, img, retain_graph=True)
    g_loss, = torch.autograd.grad(loss, img)
    out.update({name + "_img": img.detach().numpy(), name + "_gt": gt.numpy(),
                name + "_l1": np.float64(Ll1.item()), name + "_ssim": np.float64(s.item()),
                name + "_loss": np.float64(loss.item()), name + "_g_l1": g_l1.numpy(), name + "_g_ssim": g_ss.numpy(),
                name + "_g_loss": g_loss.numpy()})
np.savez_compressed(os.path.join(HERE, "loss_ssim.npz"), **out)
print("wrote loss_ssim.npz", {k: v.shape for k, v in out.items() if hasattr(v, "shape") and v.ndim})
