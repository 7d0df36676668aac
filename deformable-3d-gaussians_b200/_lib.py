"""Loader and ctypes declarations for libd3g.so (the C ABI declared in include/d3g.h).

There is deliberately no fallback: if the library is missing and cannot be compiled, importing the
product path raises.  Nothing in this package imports oracle/.
"""
import ctypes as C
import glob
import os
import shutil
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.path.join(_PKG, "libd3g.so")
CSRC = os.path.join(_PKG, "csrc")
INCLUDE = os.path.join(_ROOT, "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    deps = _sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h"))
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    """Compile every CUDA source for sm_100a into libd3g.so (in-tree, so it ships to the GPU box)."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("libd3g.so is missing/stale and nvcc was not found; cannot build the CUDA path")
    objs = []
    build_dir = os.path.join(_PKG, "build")
    os.makedirs(build_dir, exist_ok=True)
    procs = []
    for src in _sources():
        obj = os.path.join(build_dir, os.path.basename(src) + ".o")
        objs.append(obj)
        if (not force and os.path.exists(obj) and os.path.getmtime(obj) > os.path.getmtime(src)
                and all(os.path.getmtime(obj) > os.path.getmtime(h)
                        for h in glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(INCLUDE, "*.h")))):
            continue
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
               "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC, "-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, out))
    cmd = [nvcc, "-shared", "-o", LIB_PATH] + objs
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if out.returncode != 0:
        raise RuntimeError("link failed:\n" + out.stdout)
    return LIB_PATH


class RasterSettings(C.Structure):
    _fields_ = [("image_height", C.c_int32), ("image_width", C.c_int32), ("tanfovx", C.c_float),
                ("tanfovy", C.c_float), ("bg", C.c_void_p), ("scale_modifier", C.c_float),
                ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("sh_degree", C.c_int32),
                ("campos", C.c_void_p), ("prefiltered", C.c_int32), ("debug", C.c_int32),
                ("densify_mode", C.c_int32), ("binning_mode", C.c_int32)]


class AdamTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_int64), ("lr", C.c_float), ("weight_decay", C.c_float), ("step", C.c_int64)]


class MlpConfig(C.Structure):
    _fields_ = [("is_blender", C.c_int32), ("is_6dof", C.c_int32), ("heads", C.c_int32),
                ("D", C.c_int32), ("W", C.c_int32), ("multires", C.c_int32)]


_PARAM_FIELDS = (["timenet0_w", "timenet0_b", "timenet2_w", "timenet2_b"]
                 + [("linear_w", 8), ("linear_b", 8)]
                 + ["warp_w", "warp_b", "branch_w_w", "branch_w_b", "branch_v_w", "branch_v_b", "rot_w",
                    "rot_b", "scale_w", "scale_b"])


def _ptr_struct(name):
    fields = []
    for f in _PARAM_FIELDS:
        if isinstance(f, tuple):
            fields.append((f[0], C.c_void_p * f[1]))
        else:
            fields.append((f, C.c_void_p))
    return type(name, (C.Structure,), {"_fields_": fields})


MlpParams = _ptr_struct("MlpParams")
MlpGrads = _ptr_struct("MlpGrads")

# every symbol include/d3g.h declares; tests check the .so exports all of them
EXPORTS = ["d3g_last_error", "d3g_version", "d3g_prof_enable", "d3g_prof_count", "d3g_prof_name", "d3g_prof_collect", "d3g_raster_geom_bytes", "d3g_raster_img_bytes",
           "d3g_raster_bin_bytes", "d3g_raster_preprocess", "d3g_raster_render", "d3g_raster_backward",
           "d3g_raster_preprocess_split", "d3g_raster_backward_split",
           "d3g_mark_visible", "d3g_gaussians_prepare", "d3g_gaussians_prepare_backward",
           "d3g_gaussians_prepare_backward_into", "d3g_raster_read_header", "d3g_raster_header_async", "d3g_raster_read_state", "d3g_mlp_packed_bytes",
           "d3g_mlp_workspace_bytes", "d3g_mlp_workspace_bytes_times", "d3g_mlp_pack_weights", "d3g_mlp_forward", "d3g_mlp_forward_times", "d3g_mlp_backward", "d3g_mlp_backward_times", "d3g_mlp_read_status", "d3g_mlp_read_masks", "d3g_mlp_sm_reserve",
           "d3g_image_loss_workspace_bytes",
           "d3g_image_loss", "d3g_image_loss_backward", "d3g_image_loss_static", "d3g_image_loss_static_limits", "d3g_adam_step", "d3g_densification_stats", "d3g_knn_mean_dist2", "d3g_knn_workspace_bytes",
           "d3g_knn_mean_dist2_grid"]

# test-only probes, declared in include/d3g_selftest.h (not part of the product ABI)
TEST_EXPORTS = ["d3g_selftest_umma", "d3g_selftest_umma_raw", "d3g_selftest_umma_clk", "d3g_selftest_umma_f16"]

_lib = None


def load():
    """Return the loaded library; build it first if the sources are newer. Raises on failure."""
    global _lib
    if _lib is not None:
        return _lib
    if _stale():
        build_library()
    lib = C.CDLL(LIB_PATH)
    lib.d3g_last_error.restype = C.c_char_p
    lib.d3g_prof_name.restype = C.c_char_p
    for name in EXPORTS + TEST_EXPORTS:
        if not hasattr(lib, name):
            raise RuntimeError("libd3g.so does not export %s" % name)
    _lib = lib
    return lib


class D3GError(RuntimeError):
    pass


def check(code):
    if code != 0:
        msg = load().d3g_last_error().decode("utf-8", "replace")
        if code == -1:
            raise Exception(msg)  # the reference's wrapper raises bare Exception for bad argument combos
        raise D3GError("libd3g error %d: %s" % (code, msg))


def ptr(t):
    """Device pointer of a tensor (or NULL for None)."""
    return C.c_void_p(0 if t is None else t.data_ptr())
