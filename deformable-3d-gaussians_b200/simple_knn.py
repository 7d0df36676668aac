"""Mirror of the reference's `simple_knn._C` module (submodules/simple-knn/ext.cpp, spatial.cu:15-26): one function,
`distCUDA2(points) -> [P] float32`, the mean squared distance to the three nearest neighbours that
`GaussianModel.create_from_pcd` (scene/gaussian_model.py:105) turns into the initial scales.  Same argument meaning
and result as the reference; runs csrc/knn.cu through the C ABI (`d3g_knn_mean_dist2`) on the caller's current
stream.  CUDA only — there is no CPU path (the reference has none either)."""
import ctypes as C

import torch

from . import _lib


GRID_MIN_POINTS = 4096   # below this the single all-pairs launch is as fast and needs no workspace


def distCUDA2(points, method=None):
    """method: None (by size), 'pairs' (all pairs through shared memory) or 'grid' (uniform-grid search); same result."""
    if not isinstance(points, torch.Tensor) or not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor")
    if points.dim() != 2 or points.size(1) != 3:
        raise RuntimeError("distCUDA2: points must have shape [P, 3]")
    if points.dtype != torch.float32:
        raise RuntimeError("distCUDA2: points must be float32")   # the reference reads data<float>()
    pts = points.detach().contiguous()
    P = pts.size(0)
    means = torch.zeros(P, dtype=torch.float32, device=pts.device)   # reference: torch::full({P}, 0.0)
    with torch.cuda.device(pts.device):
        stream = torch.cuda.current_stream().cuda_stream
        lib = _lib.load()
        if method is None:
            method = "grid" if P >= GRID_MIN_POINTS else "pairs"
        if method == "pairs":
            _lib.check(lib.d3g_knn_mean_dist2(C.c_int32(P), _lib.ptr(pts), _lib.ptr(means), C.c_void_p(stream)))
        elif method == "grid":
            nbytes = C.c_size_t(0)
            _lib.check(lib.d3g_knn_workspace_bytes(C.c_int32(P), C.byref(nbytes)))
            ws = torch.empty(nbytes.value, dtype=torch.uint8, device=pts.device)
            _lib.check(lib.d3g_knn_mean_dist2_grid(C.c_int32(P), _lib.ptr(pts), _lib.ptr(means), _lib.ptr(ws),
                                                   C.c_size_t(nbytes.value), C.c_void_p(stream)))
        else:
            raise ValueError("distCUDA2: method must be None, 'pairs' or 'grid'")
    return means


import sys as _sys

_C = _sys.modules[__name__]   # `from simple_knn._C import distCUDA2` works when this module is registered as simple_knn
