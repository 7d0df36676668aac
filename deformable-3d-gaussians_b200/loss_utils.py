"""Image losses of the training step, same names and semantics as the reference's utils/loss_utils.py
(l1_loss :18-19, ssim :42-73), backed by ONE fused CUDA pass (csrc/loss.cu) instead of 5 grouped conv2d and ~15
elementwise launches, plus `training_loss` = the combination of train_baseline.py:126-127 / train.py:245-247.
CUDA only: CPU tensors raise."""
import ctypes as C

import torch

from . import _lib


class _ImageLoss(torch.autograd.Function):
    """value = w_l1 * mean|img - gt| + w_ssim * mean SSIM(img, gt) + bias; gradient w.r.t. img only (the reference
    never differentiates the ground truth).  Forward leaves three per-pixel partial maps in a workspace; backward turns them into the gradient, scaled by the
    upstream cotangent inside the same kernel."""

    @staticmethod
    def forward(ctx, img, gt, w_l1, w_ssim, bias):
        if img.device.type != "cuda":
            raise RuntimeError("image loss: tensors must live on a CUDA device (there is no CPU path)")
        if img.shape != gt.shape or img.dim() not in (3, 4):
            raise RuntimeError("image loss: img and gt must both be [C,H,W] or [B,C,H,W]")
        lib = _lib.load()
        x = img.detach().contiguous().float()
        y = gt.detach().contiguous().float()
        H, W = int(x.shape[-2]), int(x.shape[-1])
        Cn = int(x.numel() // (H * W))   # batch and channel planes are independent windows
        nb = C.c_size_t()
        _lib.check(lib.d3g_image_loss_workspace_bytes(Cn, H, W, C.byref(nb)))
        ws = torch.empty(nb.value, dtype=torch.uint8, device=x.device)
        out = torch.empty(3, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(lib.d3g_image_loss(Cn, H, W, _lib.ptr(x), _lib.ptr(y), C.c_float(w_l1), C.c_float(w_ssim),
                                          C.c_float(bias), _lib.ptr(ws), _lib.ptr(out), None,
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        ctx.saved = (x, y, ws, Cn, H, W, w_l1, w_ssim, img.shape)
        ctx.terms = out
        return out[2]

    @staticmethod
    def backward(ctx, g):
        if not ctx.needs_input_grad[0]:
            return None, None, None, None, None
        lib = _lib.load()
        x, y, ws, Cn, H, W, w_l1, w_ssim, shape = ctx.saved
        g = g.detach().reshape(1).float().contiguous()
        grad = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(lib.d3g_image_loss_backward(Cn, H, W, _lib.ptr(x), _lib.ptr(y), C.c_float(w_l1),
                                                   C.c_float(w_ssim), _lib.ptr(ws), _lib.ptr(g), _lib.ptr(grad),
                                                   C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return grad.view(shape), None, None, None, None


def l1_loss(network_output, gt):
    return _ImageLoss.apply(network_output, gt, 1.0, 0.0, 0.0)


def ssim(img1, img2, window_size=11, size_average=True):
    if window_size != 11 or not size_average:
        raise NotImplementedError("fused ssim implements the reference's only call form: window 11, size_average=True")
    return _ImageLoss.apply(img1, img2, 0.0, 1.0, 0.0)


def training_loss(image, gt_image, lambda_dssim=0.2):
    """(1 - lambda) * l1_loss + lambda * (1 - ssim), arguments as in train_baseline.py:126-127 (default lambda:
    arguments/__init__.py:119)."""
    lam = float(lambda_dssim)
    return _ImageLoss.apply(image, gt_image, 1.0 - lam, -lam, lam)
