"""TEST INFRASTRUCTURE — CPU restatement (numpy, float64 or float32) of the reference's image losses:
utils/loss_utils.py:18-19 (l1_loss), :31-33 (gaussian window), :36-40 (create_window), :42-73 (ssim/_ssim) and the
combination of train_baseline.py:126-127.  Pinned against the reference itself: tests/golden/loss_ssim.npz holds
values and autograd gradients produced by importing /root/reference/utils/loss_utils.py (tests/golden/make_golden.py).
Only tests/, __graft_entry__.smoke() and bench.py's CPU leg may import this."""
import math

import numpy as np


def window_1d(dtype=np.float64):
    # loss_utils.py:31-33: python-float exponentials stored as float32, normalised by their float32 sum
    g = np.array([math.exp(-(x - 11 // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)], np.float32)
    return (g / g.sum(dtype=np.float32)).astype(dtype)


def _conv(img, w):
    """zero-padded 'same' separable 11-tap filter per channel (F.conv2d(padding=5, groups=C), :57-64)."""
    C, H, W = img.shape
    p = np.zeros((C, H + 10, W + 10), img.dtype)
    p[:, 5:5 + H, 5:5 + W] = img
    hz = sum(w[k] * p[:, :, k:k + W] for k in range(11))
    return sum(w[k] * hz[:, k:k + H, :] for k in range(11))


def l1_loss(x, y):
    return np.abs(x - y).mean()


def ssim_terms(x, y):
    w = window_1d(x.dtype)
    mu1, mu2 = _conv(x, w), _conv(y, w)
    e11, e22, e12 = _conv(x * x, w), _conv(y * y, w), _conv(x * y, w)
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    s1, s2, s12 = e11 - mu1 * mu1, e22 - mu2 * mu2, e12 - mu1 * mu2
    A1, A2, B1, B2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2, mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    return w, mu1, mu2, A1, A2, B1, B2


def ssim(x, y):
    _, _, _, A1, A2, B1, B2 = ssim_terms(x, y)
    return (A1 * A2 / (B1 * B2)).mean()


def ssim_grad(x, y):
    """d ssim(x, y) / dx, analytic (chain rule through the five window moments)."""
    w, mu1, mu2, A1, A2, B1, B2 = ssim_terms(x, y)
    S = A1 * A2 / (B1 * B2)
    dmu = 2 * mu2 * (A2 - A1) / (B1 * B2) - S * 2 * mu1 * (1 / B1 - 1 / B2)
    d11 = -S / B2
    d12 = 2 * A1 / (B1 * B2)
    return (_conv(dmu, w) + 2 * x * _conv(d11, w) + y * _conv(d12, w)) / x.size


def training_loss(x, y, lam=0.2):
    return (1 - lam) * l1_loss(x, y) + lam * (1 - ssim(x, y))


def training_loss_grad(x, y, lam=0.2):
    return (1 - lam) * np.sign(x - y) / x.size - lam * ssim_grad(x, y)
