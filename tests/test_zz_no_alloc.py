"""d3g_image_loss_static (include/d3g.h): the fused L1 + SSIM loss with its workspace in a static device array, for
callers that cannot allocate.  It has no counterpart in the reference,
so the check is against the existing entry point d3g_image_loss, through the C ABI, on two calls in a row that share
the static array.  This file is named so that pytest collects it last."""
import ctypes as C

import pytest
import torch

from conftest import pkg

VALUE_TOL = 1e-5   # loss values, as tests/test_gpu_loss.py
GRAD_TOL = 1e-3    # gradients, relative to the largest reference entry (north_star: gradients <= 1e-3 relative)


def _limits(lib):
    n, b, s = C.c_size_t(), C.c_size_t(), C.c_size_t()
    assert lib.d3g_image_loss_static_limits(C.byref(n), C.byref(b), C.byref(s)) == 0
    return n.value, b.value, s.value


def test_static_loss_limits_and_oversize_rejection():
    """Shapes beyond the stated maximum are refused with a clear error before any CUDA call (runs without a GPU; the
    pointers are never dereferenced)."""
    _lib = pkg("_lib")
    lib = _lib.load()
    max_elems, max_blocks, scratch = _limits(lib)
    assert max_elems == 1 << 21 and max_blocks == 1 << 14
    assert scratch == 3 * max_elems * 4 + 2 * max_blocks * 4 and scratch < 32 << 20   # "modest": 24.1 MiB
    assert 3 * 800 * 800 <= max_elems                                              # the bench resolution fits
    fake = C.c_void_p(256)
    args = lambda c, h, w: (c, h, w, fake, fake, C.c_float(0.8), C.c_float(-0.2), C.c_float(0.2), fake, fake,
                            C.c_void_p(0))
    assert lib.d3g_image_loss_static(*args(3, 1024, 1024)) == -3            # D3G_ERR_CAPACITY: too many elements
    msg = lib.d3g_last_error().decode()
    assert "static scratch limit" in msg and str(max_elems) in msg and "d3g_image_loss" in msg
    assert lib.d3g_image_loss_static(*args(max_blocks + 1, 1, 1)) == -3      # few elements, too many pixel blocks
    assert "pixel blocks" in lib.d3g_last_error().decode()
    assert lib.d3g_image_loss_static(*args(0, 8, 8)) == -1                   # same argument checks as d3g_image_loss
    assert lib.d3g_image_loss_static(3, 8, 8, C.c_void_p(0), fake, C.c_float(1), C.c_float(0), C.c_float(0), fake,
                                     fake, C.c_void_p(0)) == -1


def _both(lib, _lib, img, gt, weights):
    """(out3, grad) of the existing entry point (caller-provided workspace) and of the static-scratch one."""
    c, h, w = img.shape
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    w_l1, w_ssim, bias = (C.c_float(v) for v in weights)
    wsb = C.c_size_t()
    _lib.check(lib.d3g_image_loss_workspace_bytes(c, h, w, C.byref(wsb)))
    ws = torch.empty(wsb.value, dtype=torch.uint8, device=img.device)
    want_out, want_grad = torch.zeros(3, device=img.device), torch.empty_like(img)
    got_out, got_grad = torch.zeros(3, device=img.device), torch.empty_like(img)
    _lib.check(lib.d3g_image_loss(c, h, w, _lib.ptr(img), _lib.ptr(gt), w_l1, w_ssim, bias, _lib.ptr(ws),
                                  _lib.ptr(want_out), _lib.ptr(want_grad), stream))
    torch.cuda.synchronize()
    # everything the call needs exists before it; the call itself takes inputs, outputs and a stream only
    _lib.check(lib.d3g_image_loss_static(c, h, w, _lib.ptr(img), _lib.ptr(gt), w_l1, w_ssim, bias,
                                         _lib.ptr(got_out), _lib.ptr(got_grad), stream))
    torch.cuda.synchronize()
    return want_out, want_grad, got_out, got_grad


@pytest.mark.gpu
def test_static_loss_matches_workspace_entry_point_on_two_calls_in_a_row():
    _lib = pkg("_lib")
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(41)
    # call 1: the bench resolution (fills most of the array); call 2: a smaller, ragged image (not a multiple of the
    # 32x16 pixel block) that finds call 1's contents in the array; call 3: call 1's shape again, other weights
    cases = [((3, 800, 800), (0.8, -0.2, 0.2)), ((3, 197, 333), (0.8, -0.2, 0.2)), ((3, 800, 800), (0.0, 1.0, 0.0))]
    with torch.cuda.device(dev):
        for shape, weights in cases:
            gt = torch.rand(shape, generator=g).to(dev)
            img = (gt + 0.1 * torch.randn(shape, generator=g).to(dev)).clamp(0, 1).contiguous()
            want_out, want_grad, got_out, got_grad = _both(lib, _lib, img, gt, weights)
            assert torch.isfinite(got_out).all() and float(want_out[0]) > 0
            assert (got_out - want_out).abs().max().item() <= VALUE_TOL, (shape, got_out, want_out)
            scale = want_grad.abs().max().item()
            assert scale > 0
            assert (got_grad - want_grad).abs().max().item() <= GRAD_TOL * scale, shape
        # value only (dL_dimg = NULL), as d3g_image_loss allows
        out = torch.zeros(3, device=dev)
        img_s, gt_s = img[:, :197, :333].contiguous(), gt[:, :197, :333].contiguous()
        _lib.check(lib.d3g_image_loss_static(3, 197, 333, _lib.ptr(img_s), _lib.ptr(gt_s), C.c_float(0.8),
                                             C.c_float(-0.2), C.c_float(0.2), _lib.ptr(out), C.c_void_p(0),
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        torch.cuda.synchronize()
        assert torch.isfinite(out).all() and 0 < float(out[0]) < 1
