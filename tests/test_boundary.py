"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol
include/d3g.h declares; the product package never touches oracle/; the Python surface mirrors the
reference's names."""
import ctypes as C
import os
import re

from conftest import ROOT, pkg


def test_library_exports_every_declared_symbol():
    _lib = pkg("_lib")
    header = open(os.path.join(ROOT, "include", "d3g.h")).read()
    declared = set(re.findall(r"\b(d3g_[a-z0-9_]+)\s*\(", header))
    declared -= {"d3g_status"}
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    lib.d3g_version.restype = C.c_int
    assert lib.d3g_version() >= 2
    # the tcgen05 probes live behind their own test-only header, outside the product ABI
    probes = open(os.path.join(ROOT, "include", "d3g_selftest.h")).read()
    declared_t = set(re.findall(r"\b(d3g_[a-z0-9_]+)\s*\(", probes))
    assert declared_t == set(_lib.TEST_EXPORTS), declared_t ^ set(_lib.TEST_EXPORTS)
    assert not declared_t & declared
    for name in declared_t:
        assert hasattr(lib, name), name


def test_size_queries_run_without_a_gpu():
    _lib = pkg("_lib")
    lib = _lib.load()
    b = C.c_size_t()
    assert lib.d3g_raster_geom_bytes(100000, C.byref(b)) == 0 and b.value > 100000 * 100
    assert lib.d3g_raster_img_bytes(800, 800, C.byref(b)) == 0 and b.value >= 800 * 800 * 8
    assert lib.d3g_raster_bin_bytes(C.c_uint64(800000), C.byref(b)) == 0 and b.value >= 800000 * 24
    cfg = _lib.MlpConfig(1, 0, 7, 8, 256, 10)
    assert lib.d3g_mlp_packed_bytes(C.byref(cfg), C.byref(b)) == 0 and b.value > 3_000_000
    assert lib.d3g_mlp_workspace_bytes(C.byref(cfg), C.c_int64(100000), 1, C.byref(b)) == 0
    bad = _lib.MlpConfig(1, 0, 7, 6, 128, 10)
    assert lib.d3g_mlp_packed_bytes(C.byref(bad), C.byref(b)) == -4
    assert b"D=8" in lib.d3g_last_error()


def test_product_never_imports_the_oracle():
    pkg_dir = os.path.join(ROOT, "deformable-3d-gaussians_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                # comments may cite the oracle; importing, including, linking or dlopen-ing it may not happen
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert not re.search(r"#\s*include\s*[\"<][^\">]*oracle", src), f
                assert "libraster_oracle" not in src and "import_module(\"oracle" not in src, f


def test_python_surface_mirrors_reference_names():
    rast = pkg("rasterizer")
    assert rast.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")
    import inspect
    renderer = pkg("renderer")
    assert list(inspect.signature(renderer.render).parameters) == [
        "viewpoint_camera", "pc", "pipe", "bg_color", "d_xyz", "d_rotation", "d_scaling", "is_6dof",
        "scaling_modifier", "override_color", "direct_compute"]
    deform = pkg("deform")
    for name in ("DeformNetwork", "DeformNetworkBaseline", "DeformModel", "DeformModelBaseline"):
        assert hasattr(deform, name)
    assert list(inspect.signature(deform.DeformNetwork.__init__).parameters)[1:] == [
        "D", "W", "input_ch", "output_ch", "multires", "is_blender", "is_6dof"]


def test_split_sh_entry_points_check_their_arguments_without_a_gpu():
    """d3g_raster_preprocess_split / d3g_raster_backward_split refuse layouts they cannot read before any CUDA call:
    only the reference's 1 + 15 coefficient split, 16-byte aligned _features_rest / its .grad."""
    _lib = pkg("_lib")
    lib = _lib.load()
    s = _lib.RasterSettings()
    s.image_height = s.image_width = 64
    s.tanfovx = s.tanfovy = 0.5
    s.bg = s.viewmatrix = s.projmatrix = s.campos = 256
    s.scale_modifier, s.sh_degree = 1.0, 3
    p, odd, null = C.c_void_p(256), C.c_void_p(260), C.c_void_p(0)
    pre = lambda dc, rest, M: lib.d3g_raster_preprocess_split(C.byref(s), 10, p, dc, rest, M, p, p, p, null, p, p,
                                                             null, null)
    assert pre(p, p, 9) == -4 and b"sh_coeffs = 16" in lib.d3g_last_error()
    assert pre(p, null, 16) == -1 and pre(null, p, 16) == -1
    assert pre(p, odd, 16) == -1 and b"16-byte aligned" in lib.d3g_last_error()
    bwd = lambda gdc, grest: lib.d3g_raster_backward_split(C.byref(s), 10, C.c_uint64(100), p, p, p, 16, p, p, null, p,
                                                           p, p, p, p, null, p, p, p, gdc, grest, p, p, p, null, null)
    assert bwd(null, p) == -1 and b"grad_features" in lib.d3g_last_error()
    assert bwd(p, odd) == -1 and b"16-byte aligned" in lib.d3g_last_error()
