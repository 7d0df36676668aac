"""Drop-in for the reference's `diff_gaussian_rasterization` Python module.

Same names, argument meaning and error behaviour as the module imported at
gaussian_renderer/__init__.py:14 and used at :53-68 / :115-124: `GaussianRasterizationSettings` is a
NamedTuple with the twelve fields the reference passes by keyword, `GaussianRasterizer(raster_settings)`
is an nn.Module whose call returns `(color[3,H,W], radii[N] int32, depth[1,H,W])`, and gradients flow
to means3D, means2D, means2D_densify, shs / colors_precomp, opacities, scales, rotations /
cov3D_precomp.  All compute goes through the C ABI of libd3g.so (include/d3g.h); this file only
forwards data pointers and the current CUDA stream.
"""
import ctypes as C
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


# Semantics of means2D_densify.grad (DESIGN.md "open semantic #1"); 2 = sum of |per-pixel term|.
DENSIFY_MODE = 2

# How the (tile, Gaussian) pairs are binned: 0 = per-tile binning (default), 1 = the reference's duplicateWithKeys +
# 64-bit radix sort + identifyTileRanges.  Both give bit-identical sorted keys, point list and tile ranges
# (tests/test_gpu_raster.py); the unsorted key list and the per-Gaussian offsets only exist in mode 1.
BINNING_MODE = 0

# Optional sync-free operation: when set to a callable (N, H, W) -> pair capacity, forward sizes the
# binning workspace from it instead of reading the pair count back from the device.
PAIR_CAPACITY_POLICY = None



class GrowingPairCapacity:
    """Sync-free capacity policy for PAIR_CAPACITY_POLICY.  The first frame of a shape reads the pair count R back
    like the reference does (`num_rendered`); later frames size the binning workspace from the largest R seen so
    far times `headroom` and only enqueue an asynchronous copy of the device header {R, overflow}.  `check()`
    (call it wherever the host synchronises anyway, e.g. after reading the loss) raises if a frame overflowed
    its capacity — the library flags that frame's image as undefined — and grows the capacity for the next one."""

    def __init__(self, headroom=1.25):
        self.headroom, self.seen, self.pending, self.free = float(headroom), {}, [], []
        self.graph_headers = []   # (key, cap, pinned header) of frames captured into CUDA graphs
        self.graph_pool = None    # pinned slots for them (pinned memory cannot be allocated during a capture)

    def __call__(self, N, H, W):
        if torch.cuda.is_current_stream_capturing():
            r = self.seen.get((N, H, W))
            if r is None:
                raise RuntimeError("GrowingPairCapacity: shape (%d,%d,%d) must be rendered once before it is captured "
                                   "into a CUDA graph" % (N, H, W))
            return int(r * self.headroom) + 1024
        self._drain(block=False)
        r = self.seen.get((N, H, W))
        return None if r is None else int(r * self.headroom) + 1024

    def observe(self, key, cap, geom, lib):
        if torch.cuda.is_current_stream_capturing():
            # the header copy becomes a node of the graph; check() examines the slot after a replay has finished
            n = len(self.graph_headers)
            if self.graph_pool is None or 2 * n + 2 > self.graph_pool.numel():
                raise RuntimeError("GrowingPairCapacity: render eagerly once before capturing into a CUDA graph")
            host = self.graph_pool[2 * n:2 * n + 2]
            _lib.check(lib.d3g_raster_header_async(_lib.ptr(geom), C.c_void_p(host.data_ptr()),
                                                   C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            self.graph_headers.append((key, cap, host))
            return
        if self.graph_pool is None:
            self.graph_pool = torch.zeros(512, dtype=torch.int32).pin_memory()
        host = self.free.pop() if self.free else torch.zeros(2, dtype=torch.int32).pin_memory()
        stream = torch.cuda.current_stream()
        _lib.check(lib.d3g_raster_header_async(_lib.ptr(geom), C.c_void_p(host.data_ptr()), C.c_void_p(stream.cuda_stream)))
        ev = torch.cuda.Event()
        ev.record(stream)
        self.pending.append((key, cap, host, ev, geom))

    def record_exact(self, key, R):
        self.seen[key] = max(self.seen.get(key, 0), R)

    def _drain(self, block):
        overflowed = None
        while self.pending:
            key, cap, host, ev, _geom = self.pending[0]
            if not block and not ev.query():
                break
            ev.synchronize()
            self.pending.pop(0)
            self.free.append(host)
            R = int(host[0].item()) & 0xFFFFFFFF
            self.seen[key] = max(self.seen.get(key, 0), R)
            if int(host[1].item()) != 0 or R > cap:
                overflowed = (key, cap, R)
        return overflowed

    def check(self):
        bad = self._drain(block=True)
        if self.graph_headers:
            torch.cuda.current_stream().synchronize()
            for key, cap, host in self.graph_headers:
                R = int(host[0].item()) & 0xFFFFFFFF
                if int(host[1].item()) != 0 or R > cap:
                    raise RuntimeError("GaussianRasterizer: %d pairs exceeded the capacity %d baked into a CUDA graph "
                                       "(N,H,W=%s); re-capture it with more headroom" % (R, cap, key))
        if bad is not None:
            raise RuntimeError("GaussianRasterizer: %d pairs exceeded the capacity %d of a sync-free frame (N,H,W=%s); "
                               "that frame's image is undefined — re-render it" % (bad[2], bad[1], bad[0]))


_pinned = {}


def _pinned_u32(device):
    key = (device.index if device.index is not None else torch.cuda.current_device())
    if key not in _pinned:
        _pinned[key] = torch.zeros(4, dtype=torch.int32).pin_memory()
    return _pinned[key]


def _f32c(t):
    if t is None:
        return None
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.contiguous().float()
    return t


def _settings_struct(rs, keep):
    """Build the C struct; `keep` collects the tensors whose pointers it holds."""
    bg, view, proj, campos = (_f32c(rs.bg), _f32c(rs.viewmatrix), _f32c(rs.projmatrix), _f32c(rs.campos))
    keep.extend([bg, view, proj, campos])
    s = _lib.RasterSettings()
    s.image_height, s.image_width = int(rs.image_height), int(rs.image_width)
    s.tanfovx, s.tanfovy = float(rs.tanfovx), float(rs.tanfovy)
    s.bg, s.viewmatrix, s.projmatrix, s.campos = bg.data_ptr(), view.data_ptr(), proj.data_ptr(), campos.data_ptr()
    s.scale_modifier = float(rs.scale_modifier)
    s.sh_degree = int(rs.sh_degree)
    s.prefiltered, s.debug = int(bool(rs.prefiltered)), int(bool(rs.debug))
    s.densify_mode = DENSIFY_MODE
    s.binning_mode = BINNING_MODE
    return s


def _sizes(lib, N, H, W):
    gb, ib = C.c_size_t(), C.c_size_t()
    _lib.check(lib.d3g_raster_geom_bytes(N, C.byref(gb)))
    _lib.check(lib.d3g_raster_img_bytes(H, W, C.byref(ib)))
    return gb.value, ib.value


class RasterState:
    """Opaque forward state (geometry / binning / image workspaces) kept for backward and for parity
    inspection through d3g_raster_read_state."""

    def __init__(self, N, H, W, cap, geom, binb, img, R):
        self.N, self.H, self.W, self.cap, self.geom, self.bin, self.img, self.R = N, H, W, cap, geom, binb, img, R

    def read(self, which, dtype, count):
        lib = _lib.load()
        out = torch.empty(count, dtype=dtype)
        stream = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.d3g_raster_read_state(which, self.N, self.H, self.W, C.c_uint64(self.cap),
                                             _lib.ptr(self.geom), _lib.ptr(self.bin), _lib.ptr(self.img),
                                             C.c_void_p(out.data_ptr()), C.c_size_t(out.numel() * out.element_size()),
                                             C.c_void_p(stream)))
        return out


def rasterize_forward(rs, means3D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp, sh_split=None):
    """Forward through the C ABI. Returns (color, radii, depth, RasterState).  sh_split = (features_dc [N,1,3],
    features_rest [N,15,3]) replaces `shs` by the two tensors it would be concatenated from."""
    dev = means3D.device
    if dev.type != "cuda":
        raise RuntimeError("GaussianRasterizer: tensors must live on a CUDA device (there is no CPU path)")
    with torch.cuda.device(dev):
        lib = _lib.load()
        keep = []
        s = _settings_struct(rs, keep)
        N = int(means3D.shape[0])
        H, W = s.image_height, s.image_width
        M = int(shs.shape[1]) if shs is not None else 0
        stream = torch.cuda.current_stream().cuda_stream
        gb, ib = _sizes(lib, N, H, W)
        geom = torch.empty(gb, dtype=torch.uint8, device=dev)
        img = torch.empty(ib, dtype=torch.uint8, device=dev)
        radii = torch.empty(N, dtype=torch.int32, device=dev)
        color = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        depth = torch.empty(1, H, W, dtype=torch.float32, device=dev)
        policy = PAIR_CAPACITY_POLICY
        host = None if policy is not None else _pinned_u32(dev)
        if sh_split is not None:
            _lib.check(lib.d3g_raster_preprocess_split(
                C.byref(s), N, _lib.ptr(means3D), _lib.ptr(sh_split[0]), _lib.ptr(sh_split[1]),
                int(sh_split[0].shape[1] + sh_split[1].shape[1]), _lib.ptr(opacities), _lib.ptr(scales),
                _lib.ptr(rotations), _lib.ptr(cov3D_precomp), _lib.ptr(geom), _lib.ptr(radii),
                C.c_void_p(0 if host is None else host.data_ptr()), C.c_void_p(stream)))
        else:
            _lib.check(lib.d3g_raster_preprocess(C.byref(s), N, _lib.ptr(means3D), _lib.ptr(shs), M,
                                                 _lib.ptr(colors_precomp), _lib.ptr(opacities), _lib.ptr(scales),
                                                 _lib.ptr(rotations), _lib.ptr(cov3D_precomp), _lib.ptr(geom),
                                                 _lib.ptr(radii), C.c_void_p(0 if host is None else host.data_ptr()),
                                                 C.c_void_p(stream)))
        cap = None if policy is None else policy(N, H, W)
        if cap is None:
            if host is None:
                host = _pinned_u32(dev)
                _lib.check(lib.d3g_raster_header_async(_lib.ptr(geom), C.c_void_p(host.data_ptr()), C.c_void_p(stream)))
            torch.cuda.current_stream().synchronize()
            R = int(host[0].item()) & 0xFFFFFFFF
            cap = R
            if hasattr(policy, "record_exact"):
                policy.record_exact((N, H, W), R)
        else:
            R = None
            cap = int(cap)
        bb = C.c_size_t()
        _lib.check(lib.d3g_raster_bin_bytes(C.c_uint64(cap), C.byref(bb)))
        binb = torch.empty(bb.value, dtype=torch.uint8, device=dev)
        _lib.check(lib.d3g_raster_render(C.byref(s), N, C.c_uint64(cap), _lib.ptr(geom), _lib.ptr(binb),
                                         _lib.ptr(img), _lib.ptr(radii), _lib.ptr(color), _lib.ptr(depth),
                                         C.c_void_p(stream)))
        if R is None and hasattr(policy, "observe"):
            policy.observe((N, H, W), cap, geom, lib)   # {R, overflow} of this frame, checked off the critical path
        return color, radii, depth, RasterState(N, H, W, cap, geom, binb, img, R)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, means2D_densify, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings, sh_split=None):
        # sh_split = (features_dc, features_rest, grad_dc, grad_rest): the SH coefficients are read from the two
        # parameters in place and their gradients are ADDED to the two .grad tensors by backward (render() sets
        # this up under renderer.FUSED_GRAD_ACCUMULATION); `sh` and `colors_precomp` are then None
        means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp = map(
            _f32c, (means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp))
        color, radii, depth, state = rasterize_forward(raster_settings, means3D, sh, colors_precomp, opacities,
                                                       scales, rotations, cov3Ds_precomp,
                                                       None if sh_split is None else sh_split[:2])
        ctx.raster_settings = raster_settings
        ctx.state = state
        ctx.sh_split = sh_split
        ctx.has = (sh is not None, colors_precomp is not None, scales is not None, cov3Ds_precomp is not None)
        ctx.save_for_backward(means3D, sh, colors_precomp, scales, rotations, cov3Ds_precomp, radii)
        ctx.mark_non_differentiable(radii)
        ctx.set_materialize_grads(False)
        _RasterizeGaussians.last_state = state
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_color, _grad_radii, grad_depth):
        with torch.cuda.device(ctx.saved_tensors[0].device):
            return _RasterizeGaussians._backward(ctx, grad_color, _grad_radii, grad_depth)

    @staticmethod
    def _backward(ctx, grad_color, _grad_radii, grad_depth):
        lib = _lib.load()
        means3D, sh, colors_precomp, scales, rotations, cov3Ds_precomp, radii = ctx.saved_tensors
        st = ctx.state
        dev = means3D.device
        N, H, W = st.N, st.H, st.W
        keep = []
        s = _settings_struct(ctx.raster_settings, keep)
        grad_color = torch.zeros(3, H, W, device=dev) if grad_color is None else _f32c(grad_color)
        grad_depth = _f32c(grad_depth)
        M = int(sh.shape[1]) if sh is not None else 0
        new = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
        g_means3D, g_means2D, g_densify, g_opac = new(N, 3), new(N, 3), new(N, 3), new(N, 1)
        g_sh = new(N, M, 3) if sh is not None else None
        g_colors = new(N, 3) if colors_precomp is not None else None
        g_scales = new(N, 3) if scales is not None else None
        g_rots = new(N, 4) if rotations is not None else None
        g_cov = new(N, 6) if cov3Ds_precomp is not None else None
        stream = torch.cuda.current_stream().cuda_stream
        if ctx.sh_split is not None:
            dc, rest, grad_dc, grad_rest = ctx.sh_split
            _lib.check(lib.d3g_raster_backward_split(
                C.byref(s), N, C.c_uint64(st.cap), _lib.ptr(means3D), _lib.ptr(dc), _lib.ptr(rest),
                int(dc.shape[1] + rest.shape[1]), _lib.ptr(scales), _lib.ptr(rotations), _lib.ptr(cov3Ds_precomp),
                _lib.ptr(radii), _lib.ptr(st.geom), _lib.ptr(st.bin), _lib.ptr(st.img), _lib.ptr(grad_color),
                _lib.ptr(grad_depth), _lib.ptr(g_means3D), _lib.ptr(g_means2D), _lib.ptr(g_densify),
                _lib.ptr(grad_dc), _lib.ptr(grad_rest), _lib.ptr(g_opac), _lib.ptr(g_scales), _lib.ptr(g_rots),
                _lib.ptr(g_cov), C.c_void_p(stream)))
            return (g_means3D, g_means2D, g_densify, None, None, g_opac, g_scales, g_rots, g_cov, None, None)
        _lib.check(lib.d3g_raster_backward(
            C.byref(s), N, C.c_uint64(st.cap), _lib.ptr(means3D), _lib.ptr(sh), M, _lib.ptr(colors_precomp),
            _lib.ptr(scales), _lib.ptr(rotations), _lib.ptr(cov3Ds_precomp), _lib.ptr(radii), _lib.ptr(st.geom),
            _lib.ptr(st.bin), _lib.ptr(st.img), _lib.ptr(grad_color), _lib.ptr(grad_depth), _lib.ptr(g_means3D),
            _lib.ptr(g_means2D), _lib.ptr(g_densify), _lib.ptr(g_sh), _lib.ptr(g_colors), _lib.ptr(g_opac),
            _lib.ptr(g_scales), _lib.ptr(g_rots), _lib.ptr(g_cov), C.c_void_p(stream)))
        return (g_means3D, g_means2D, g_densify, g_sh, g_colors, g_opac, g_scales, g_rots, g_cov, None, None)


def rasterize_gaussians(means3D, means2D, means2D_densify, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings, sh_split=None):
    return _RasterizeGaussians.apply(means3D, means2D, means2D_densify, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings, sh_split)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        lib = _lib.load()
        with torch.no_grad(), torch.cuda.device(positions.device):
            positions = _f32c(positions)
            view = _f32c(self.raster_settings.viewmatrix)
            out = torch.empty(positions.shape[0], dtype=torch.uint8, device=positions.device)
            _lib.check(lib.d3g_mark_visible(int(positions.shape[0]), _lib.ptr(positions), _lib.ptr(view),
                                            _lib.ptr(out), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out.bool()

    def forward(self, means3D, means2D, means2D_densify, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        return rasterize_gaussians(means3D, means2D, means2D_densify, shs, colors_precomp, opacities, scales,
                                   rotations, cov3D_precomp, self.raster_settings)
