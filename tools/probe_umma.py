"""Empirical probe of the tcgen05 shared-memory operand conventions (run on the GPU box).

For a given (major, layout type, LBO, SBO) it dumps which 32-bit word of the A (or B) image the tensor
core reads for logical element (mn, k) of one kind::tf32 MMA (K = 8), by multiplying with an identity.
"""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
_lib = importlib.import_module("deformable-3d-gaussians_b200._lib")
lib = _lib.load()


def desc(lbo, sbo, layout, addr=0, version=1):
    return (addr >> 4) | ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (version << 46) | (layout << 61)


def idesc(M, N, a_mn, b_mn):
    return (1 << 4) | (2 << 7) | (2 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def kmajor_img(X):
    """X [rows, 8] -> K-major SW128 image (rows*128 bytes, only k<8 used)."""
    rows = X.shape[0]
    img = np.zeros(rows * 32, np.float32)
    for r in range(rows):
        for k in range(8):
            img[r * 32 + (((k // 4) ^ (r & 7)) * 4) + (k % 4)] = X[r, k]
    return img


def run(a_img, b_img, ad, bd, idsc, N):
    a = torch.tensor(a_img).cuda()
    b = torch.tensor(b_img).cuda()
    D = torch.zeros(128, N, device="cuda")
    adt = torch.tensor([ad], dtype=torch.int64).cuda()
    bdt = torch.tensor([bd], dtype=torch.int64).cuda()
    _lib.check(lib.d3g_selftest_umma_raw(_lib.ptr(a), a.numel() * 4, _lib.ptr(b), b.numel() * 4, 1, _lib.ptr(adt),
                                         _lib.ptr(bdt), C.c_uint32(idsc), N, _lib.ptr(D), 1, C.c_void_p(0)))
    torch.cuda.synchronize()
    return D.cpu().numpy()


def probe_A(a_mn, layout, lbo, sbo, words=4096):
    """returns idx[m, k] = word index of the A image read for A(m,k)"""
    B = np.zeros((16, 8), np.float32)
    for k in range(8):
        B[k, k] = 1.0
    b_img = kmajor_img(B)
    res = np.zeros((128, 8), np.int64)
    for part in range(2):
        w = np.arange(words)
        vals = (w % 2048 if part == 0 else w // 2048).astype(np.float32)
        D = run(vals, b_img, desc(lbo, sbo, layout), desc(16, 1024, 2), idesc(128, 16, a_mn, 0), 16)
        res += (D[:, :8].round().astype(np.int64)) * (1 if part == 0 else 2048)
    return res


def probe_B(b_mn, layout, lbo, sbo, N=64, words=4096):
    A = np.zeros((128, 8), np.float32)
    for k in range(8):
        A[k, k] = 1.0
    a_img = kmajor_img(A)
    res = np.zeros((N, 8), np.int64)
    for part in range(2):
        w = np.arange(words)
        vals = (w % 2048 if part == 0 else w // 2048).astype(np.float32)
        D = run(a_img, vals, desc(16, 1024, 2), desc(lbo, sbo, layout), idesc(128, N, 0, b_mn), N)
        res += (D[:8, :].T.round().astype(np.int64)) * (1 if part == 0 else 2048)
    return res


def show(name, idx, rows=(0, 1, 2, 3, 4, 5, 7, 8, 9, 16, 31, 32, 33, 64, 96, 127)):
    print("==", name)
    for m in rows:
        if m < idx.shape[0]:
            print("  mn=%3d:" % m, " ".join("%5d" % (v * 4) for v in idx[m]), " (byte offsets, k=0..7)")


if __name__ == "__main__":
    show("A K-major SW128 lbo16 sbo1024 (known good)", probe_A(0, 2, 16, 1024))
    for layout in (2, 0, 1):
        for lbo, sbo in ((2048, 1024), (1024, 2048), (4096, 256)):
            try:
                show("A MN-major layout=%d lbo=%d sbo=%d" % (layout, lbo, sbo), probe_A(1, layout, lbo, sbo))
            except Exception as e:
                print("A MN layout", layout, lbo, sbo, "failed:", e)
    for layout in (2, 0):
        for lbo, sbo in ((2048, 1024), (1024, 2048)):
            try:
                show("B MN-major layout=%d lbo=%d sbo=%d" % (layout, lbo, sbo), probe_B(1, layout, lbo, sbo),
                     rows=(0, 1, 2, 3, 4, 7, 8, 31, 32, 33, 63))
            except Exception as e:
                print("B MN layout", layout, lbo, sbo, "failed:", e)
