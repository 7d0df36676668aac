"""ORACLE / TEST INFRASTRUCTURE ONLY.  ctypes front end of oracle/knn_ref.c (CPU restatement of distCUDA2) and, on a GPU
box, of oracle/_ref/libsimple_knn_ref.so — the UNMODIFIED reference `SimpleKNN::knn`
(submodules/simple-knn/simple_knn.cu) compiled by `make -C oracle ref` in the build container."""
import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_DIR, "libknn_oracle.so")
REF_SO = os.path.join(_DIR, "_ref", "libsimple_knn_ref.so")
_lib = None


def build():
    src = os.path.join(_DIR, "knn_ref.c")
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _DIR, "libknn_oracle.so"], stdout=subprocess.DEVNULL)
    return _SO


def build_ref():
    """Compile the reference's own simple_knn.cu into oracle/_ref/ (only where /root/reference exists)."""
    if os.path.isdir("/root/reference/submodules/simple-knn") and not os.path.exists(REF_SO):
        subprocess.check_call(["make", "-C", _DIR, "ref"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return REF_SO if os.path.exists(REF_SO) else None


def mean_dist2(points):
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    pts = np.ascontiguousarray(points, np.float32)
    assert pts.ndim == 2 and pts.shape[1] == 3
    out = np.zeros(pts.shape[0], np.float32)
    _lib.knn_mean_dist2_ref(C.c_int32(pts.shape[0]), pts.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    return out


def reference_mean_dist2(points_cuda):
    """Run the compiled reference on a CUDA tensor [P,3] fp32 (GPU box only). Returns a CUDA tensor [P]."""
    import torch
    lib = C.CDLL(REF_SO)
    pts = points_cuda.contiguous()
    out = torch.zeros(pts.size(0), dtype=torch.float32, device=pts.device)
    torch.cuda.synchronize()
    rc = lib.ref_simple_knn(C.c_int(pts.size(0)), C.c_void_p(pts.data_ptr()), C.c_void_p(out.data_ptr()))
    assert rc == 0, "reference knn: cuda error %d" % rc
    return out
