"""GPU: fused L1 + SSIM loss (csrc/loss.cu through the C ABI) against the reference's own values and autograd
gradients (golden fixture) and, at the bench resolution, against the oracle."""
import os

import numpy as np
import pytest
import torch

from conftest import pkg
from oracle import loss_ref

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "loss_ssim.npz"))


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["a", "b", "c"])
def test_fused_loss_matches_reference(case):
    lu = pkg("loss_utils")
    dev = torch.device("cuda", 0)
    gt = torch.tensor(G[case + "_gt"], device=dev)
    for fn, val, grad, tol in ((lambda x: lu.l1_loss(x, gt), "_l1", "_g_l1", 1e-6),
                               (lambda x: lu.ssim(x, gt), "_ssim", "_g_ssim", 1e-5),
                               (lambda x: lu.training_loss(x, gt, 0.2), "_loss", "_g_loss", 1e-5)):
        img = torch.tensor(G[case + "_img"], device=dev, requires_grad=True)
        out = fn(img)
        (2.0 * out).backward()   # exercises the upstream-cotangent scaling
        assert abs(out.item() - float(G[case + val])) < tol, (val, out.item(), float(G[case + val]))
        ref = 2.0 * G[case + grad]
        err = np.abs(img.grad.cpu().numpy() - ref).max()
        assert err <= 1e-3 * np.abs(ref).max(), (grad, err, np.abs(ref).max())   # north_star: gradients <= 1e-3 rel


@pytest.mark.gpu
def test_fused_loss_full_size_and_batched():
    lu = pkg("loss_utils")
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(3)
    gt = torch.rand(3, 800, 800, generator=g)
    img = (gt + 0.1 * torch.randn(3, 800, 800, generator=g)).clamp(0, 1)
    x = img.to(dev).requires_grad_(True)
    loss = lu.training_loss(x, gt.to(dev), 0.2)
    loss.backward()
    xn, yn = img.numpy().astype(np.float64), gt.numpy().astype(np.float64)
    assert abs(loss.item() - loss_ref.training_loss(xn, yn, 0.2)) < 1e-5
    ref = loss_ref.training_loss_grad(xn, yn, 0.2)
    assert np.abs(x.grad.cpu().numpy() - ref).max() <= 1e-3 * np.abs(ref).max()
    # identical images: ssim = 1, l1 = 0, zero gradient
    same = lu.training_loss(gt.to(dev).requires_grad_(True), gt.to(dev), 0.2)
    assert abs(same.item()) < 1e-6
    # [B,C,H,W] treats every plane independently, as the grouped conv does
    xb = torch.stack([img, gt]).to(dev)
    yb = torch.stack([gt, gt]).to(dev)
    sb = lu.ssim(xb, yb).item()
    s0 = lu.ssim(img.to(dev), gt.to(dev)).item()
    assert abs(sb - 0.5 * (s0 + 1.0)) < 1e-5
    with pytest.raises(RuntimeError):
        lu.l1_loss(img, gt)   # CPU tensors: no fallback
