"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes/numpy binding of oracle/raster_ref.c.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product package never does (tests/test_boundary.py greps for it).

The functions mirror the stages of the rasterizer the reference invokes at
gaussian_renderer/__init__.py:115-124 (see raster_ref_impl.h for the provenance statement;
PARITY UNPINNED against the reference's own CUDA extension, whose source is absent).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "libraster_oracle.so")
    src = [os.path.join(_HERE, f) for f in ("raster_ref.c", "raster_ref_impl.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.scan_tiles_f32.restype = C.c_uint64
        _LIB.scan_tiles_f64.restype = C.c_uint64
        _LIB.frame_fwd_bwd_f32.restype = C.c_uint64
        _LIB.sort_end_bit.restype = C.c_int
    return _LIB


def _cfg_struct(real):
    class Cfg(C.Structure):
        _fields_ = [("H", C.c_int32), ("W", C.c_int32), ("tanfovx", real), ("tanfovy", real),
                    ("bg", real * 3), ("scale_modifier", real), ("view", real * 16),
                    ("proj", real * 16), ("campos", real * 3), ("sh_degree", C.c_int32),
                    ("M", C.c_int32), ("densify_mode", C.c_int32)]
    return Cfg


CfgF32 = _cfg_struct(C.c_float)
CfgF64 = _cfg_struct(C.c_double)


def make_cfg(H, W, tanfovx, tanfovy, bg, scale_modifier, view, proj, campos, sh_degree, M,
             densify_mode=2, dtype=np.float32):
    """view / proj are the 4x4 arrays exactly as the reference stores them (transposed,
    scene/cameras.py:55-60); they are flattened row-major."""
    cls = CfgF32 if dtype == np.float32 else CfgF64
    c = cls()
    c.H, c.W = int(H), int(W)
    c.tanfovx, c.tanfovy = float(dtype(tanfovx)), float(dtype(tanfovy))
    for i in range(3):
        c.bg[i] = float(bg[i])
        c.campos[i] = float(campos[i])
    c.scale_modifier = float(scale_modifier)
    v = np.asarray(view, dtype=dtype).reshape(16)
    p = np.asarray(proj, dtype=dtype).reshape(16)
    for i in range(16):
        c.view[i] = float(v[i])
        c.proj[i] = float(p[i])
    c.sh_degree, c.M, c.densify_mode = int(sh_degree), int(M), int(densify_mode)
    return c


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _sfx(dtype):
    return "_f32" if dtype == np.float32 else "_f64"


def forward(cfg, means3D, shs, colors_precomp, opacities, scales, rots, cov3D_precomp,
            dtype=np.float32):
    """Full forward.  Returns a dict with every intermediate artefact."""
    L = lib()
    s = _sfx(dtype)
    cv = lambda a: None if a is None else np.ascontiguousarray(a, dtype=dtype)
    means3D, shs, colors_precomp, opacities, scales, rots, cov3D_precomp = map(
        cv, (means3D, shs, colors_precomp, opacities, scales, rots, cov3D_precomp))
    N = means3D.shape[0]
    H, W = cfg.H, cfg.W
    gx, gy = (W + 15) // 16, (H + 15) // 16
    o = dict(N=N)
    o["radii"] = np.zeros(N, np.int32)
    o["means2D"] = np.zeros((N, 2), dtype)
    o["depths"] = np.zeros(N, dtype)
    o["cov3D"] = np.zeros((N, 6), dtype)
    o["rgb"] = np.zeros((N, 3), dtype)
    o["conic_opacity"] = np.zeros((N, 4), dtype)
    o["tiles_touched"] = np.zeros(N, np.uint32)
    o["clamped"] = np.zeros((N, 3), np.uint8)
    getattr(L, "preprocess_fwd" + s)(C.byref(cfg), N, _p(means3D), _p(shs), _p(colors_precomp),
                                     _p(opacities), _p(scales), _p(rots), _p(cov3D_precomp),
                                     _p(o["radii"]), _p(o["means2D"]), _p(o["depths"]), _p(o["cov3D"]),
                                     _p(o["rgb"]), _p(o["conic_opacity"]), _p(o["tiles_touched"]),
                                     _p(o["clamped"]))
    o["offsets"] = np.zeros(N, np.uint32)
    R = int(getattr(L, "scan_tiles" + s)(N, _p(o["tiles_touched"]), _p(o["offsets"])))
    o["R"] = R
    o["keys_unsorted"] = np.zeros(max(R, 1), np.uint64)
    o["vals_unsorted"] = np.zeros(max(R, 1), np.uint32)
    getattr(L, "duplicate_with_keys" + s)(C.byref(cfg), N, _p(o["means2D"]), _p(o["depths"]),
                                          _p(o["offsets"]), _p(o["radii"]), _p(o["keys_unsorted"]),
                                          _p(o["vals_unsorted"]))
    o["keys"] = o["keys_unsorted"].copy()
    o["point_list"] = o["vals_unsorted"].copy()
    L.sort_pairs_u64(C.c_uint64(R), L.sort_end_bit(gx * gy), _p(o["keys"]), _p(o["point_list"]))
    o["ranges"] = np.zeros((gx * gy, 2), np.uint32)
    getattr(L, "identify_tile_ranges" + s)(C.c_uint64(R), _p(o["keys"]), gx * gy, _p(o["ranges"]))
    o["color"] = np.zeros((3, H, W), dtype)
    o["depth"] = np.zeros((1, H, W), dtype)
    o["final_T"] = np.zeros((H, W), dtype)
    o["n_contrib"] = np.zeros((H, W), np.uint32)
    getattr(L, "render_fwd" + s)(C.byref(cfg), _p(o["ranges"]), _p(o["point_list"]), _p(o["means2D"]),
                                 _p(o["rgb"]), _p(o["depths"]), _p(o["conic_opacity"]), _p(o["color"]),
                                 _p(o["depth"]), _p(o["final_T"]), _p(o["n_contrib"]))
    o["_inputs"] = (means3D, shs, colors_precomp, opacities, scales, rots, cov3D_precomp)
    return o


def backward(cfg, fwd, dL_dcolor_img, dL_ddepth_img=None, dtype=np.float32):
    """Full backward from image cotangents.  `fwd` is the dict returned by forward()."""
    L = lib()
    s = _sfx(dtype)
    means3D, shs, colors_precomp, opacities, scales, rots, cov3D_precomp = fwd["_inputs"]
    N = fwd["N"]
    M = cfg.M
    g = {}
    dpix = np.ascontiguousarray(dL_dcolor_img, dtype=dtype)
    ddep = None if dL_ddepth_img is None else np.ascontiguousarray(dL_ddepth_img, dtype=dtype)
    g["dL_dmean2D"] = np.zeros((N, 2), dtype)
    g["dL_dmean2D_densify"] = np.zeros((N, 2), dtype)
    g["dL_dconic"] = np.zeros((N, 3), dtype)
    g["dL_dopacity"] = np.zeros((N, 1), dtype)
    g["dL_dcolor"] = np.zeros((N, 3), dtype)
    g["dL_dz"] = np.zeros(N, dtype)
    getattr(L, "render_bwd" + s)(C.byref(cfg), _p(fwd["ranges"]), _p(fwd["point_list"]),
                                 _p(fwd["means2D"]), _p(fwd["conic_opacity"]), _p(fwd["rgb"]),
                                 _p(fwd["depths"]), _p(fwd["final_T"]), _p(fwd["n_contrib"]), _p(dpix),
                                 _p(ddep), _p(g["dL_dmean2D"]), _p(g["dL_dmean2D_densify"]),
                                 _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolor"]),
                                 _p(g["dL_dz"]))
    g["dL_dmeans3D"] = np.zeros((N, 3), dtype)
    g["dL_dsh"] = np.zeros((N, M, 3), dtype) if shs is not None else None
    g["dL_dcolors_precomp"] = np.zeros((N, 3), dtype) if colors_precomp is not None else None
    g["dL_dscales"] = np.zeros((N, 3), dtype)
    g["dL_drots"] = np.zeros((N, 4), dtype)
    g["dL_dcov3D"] = np.zeros((N, 6), dtype)
    getattr(L, "preprocess_bwd" + s)(C.byref(cfg), N, _p(means3D), _p(shs), _p(colors_precomp),
                                     _p(scales), _p(rots), _p(cov3D_precomp), _p(fwd["radii"]),
                                     _p(fwd["cov3D"]), _p(fwd["clamped"]), _p(g["dL_dmean2D"]),
                                     _p(g["dL_dconic"]), _p(g["dL_dcolor"]), _p(g["dL_dz"]),
                                     _p(g["dL_dmeans3D"]), _p(g["dL_dsh"]), _p(g["dL_dcolors_precomp"]),
                                     _p(g["dL_dscales"]), _p(g["dL_drots"]), _p(g["dL_dcov3D"]))
    return g


def frame_fwd_bwd_f32(cfg, means3D, shs, opacities, scales, rots, dL_dpix, do_backward=True):
    """Whole-frame CPU pass used for the restated CPU baseline timing. Returns (R, color, depth, grads)."""
    L = lib()
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    means3D, shs, opacities, scales, rots, dL_dpix = map(f, (means3D, shs, opacities, scales, rots, dL_dpix))
    N = means3D.shape[0]
    H, W = cfg.H, cfg.W
    color = np.zeros((3, H, W), np.float32)
    depth = np.zeros((1, H, W), np.float32)
    gm = np.zeros((N, 3), np.float32)
    gsh = np.zeros((N, cfg.M, 3), np.float32)
    gop = np.zeros((N, 1), np.float32)
    gsc = np.zeros((N, 3), np.float32)
    gro = np.zeros((N, 4), np.float32)
    R = L.frame_fwd_bwd_f32(C.byref(cfg), N, _p(means3D), _p(shs), _p(opacities), _p(scales), _p(rots),
                            _p(dL_dpix), _p(color), _p(depth), _p(gm), _p(gsh), _p(gop), _p(gsc), _p(gro),
                            C.c_uint64(0), int(do_backward))
    return int(R), color, depth, dict(means3D=gm, shs=gsh, opacities=gop, scales=gsc, rots=gro)
