"""Drop-in for `gaussian_renderer.render` (gaussian_renderer/__init__.py:32-133).

Same signature, same branching (including the quirks listed in SURVEY.md §7: un-normalised
`get_rotation + d_rotation`, deltas used verbatim under `direct_compute`, unreachable `override_color`)
and the same returned dict.  The rasterizer behind it is this package's C-ABI one.
"""
import ctypes as C
import math

import torch

from . import _lib
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians

# render() replaces the per-frame elementwise torch launches (parameter activations + deformation deltas)
# by one fused kernel each way when `pc` exposes the reference GaussianModel's raw parameters and its
# stock activation functions (scene/gaussian_model.py:58-78).  Set to False to force the torch-op path.
FUSED_PREPARE = True

# Opt-in, like DeformNetwork.fused_grad_accumulation: with the fused prepare path and a `pc` whose six raw parameters
# (_xyz, _features_dc, _features_rest, _scaling, _rotation, _opacity) already hold contiguous fp32 .grad tensors,
# backward ADDS their gradients straight into those tensors — the kernels that produce them do the read-add-write —
# and autograd receives None for them: no per-parameter AccumulateGrad add launch (six per frame), and the SH
# coefficients are read from _features_dc / _features_rest in place instead of through `pc.get_features`' 19 MB
# concatenation (and its split + strided adds in backward).  Only valid with loss.backward()-style accumulation
# into .grad (not torch.autograd.grad), exactly like the MLP flag.
FUSED_GRAD_ACCUMULATION = False
# Optional callable, invoked (with no arguments) from backward each time ONE frame's Gaussian-parameter gradients
# have all been added under FUSED_GRAD_ACCUMULATION — no post-accumulate-grad hooks fire on that path
# (frames.FlatGrads uses it to start its early all-reduce).
ON_GAUSSIAN_GRADS = None


class _PrepareGaussians(torch.autograd.Function):
    """(xyz, d_xyz, _scaling, d_scaling, _rotation, d_rotation, _opacity) -> (means3D, opacity, scales,
    rotations) exactly as gaussian_renderer/__init__.py:69-97 computes them, through d3g_gaussians_prepare."""

    @staticmethod
    def forward(ctx, xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, is_6dof, sinks=None):
        # sinks = (xyz.grad, scaling.grad, rotation.grad, opacity.grad) under FUSED_GRAD_ACCUMULATION: backward adds
        # the four parameter gradients into them and hands autograd None
        with torch.cuda.device(xyz.device):
            out = _PrepareGaussians._forward(ctx, xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity,
                                             is_6dof)
        ctx.sinks = sinks
        return out

    @staticmethod
    def _forward(ctx, xyz, d_xyz, scaling, d_scaling, rotation, d_rotation, opacity, is_6dof):
        lib = _lib.load()
        N = int(xyz.shape[0])
        dev = xyz.device
        new = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
        means3D, opac, scales, rots = new(N, 3), new(N, 1), new(N, 3), new(N, 4)
        stream = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.d3g_gaussians_prepare(N, int(is_6dof), _lib.ptr(xyz), _lib.ptr(d_xyz), _lib.ptr(scaling),
                                             _lib.ptr(d_scaling), _lib.ptr(rotation), _lib.ptr(d_rotation),
                                             _lib.ptr(opacity), _lib.ptr(means3D), _lib.ptr(opac),
                                             _lib.ptr(scales), _lib.ptr(rots), C.c_void_p(stream)))
        ctx.is_6dof = bool(is_6dof)
        ctx.save_for_backward(xyz, d_xyz, scaling, rotation, opacity)
        ctx.set_materialize_grads(False)
        return means3D, opac, scales, rots

    @staticmethod
    def backward(ctx, g_means, g_opac, g_scales, g_rots):
        with torch.cuda.device(ctx.saved_tensors[0].device):
            return _PrepareGaussians._backward(ctx, g_means, g_opac, g_scales, g_rots)

    @staticmethod
    def _backward(ctx, g_means, g_opac, g_scales, g_rots):
        lib = _lib.load()
        xyz, d_xyz, scaling, rotation, opacity = ctx.saved_tensors
        N = int(xyz.shape[0])
        dev = xyz.device
        f = lambda g: None if g is None else g.contiguous().float()
        g_means, g_opac, g_scales, g_rots = f(g_means), f(g_opac), f(g_scales), f(g_rots)
        need = ctx.needs_input_grad
        new = lambda *shape: torch.empty(*shape, dtype=torch.float32, device=dev)
        six = ctx.is_6dof
        if ctx.sinks is not None:
            acc_xyz, acc_scaling, acc_rotation, acc_opacity = ctx.sinks
            g_T = new(N, 4, 4) if (six and need[1]) else None
            stream = torch.cuda.current_stream().cuda_stream
            _lib.check(lib.d3g_gaussians_prepare_backward_into(
                N, int(six), _lib.ptr(xyz), _lib.ptr(d_xyz), _lib.ptr(scaling), _lib.ptr(rotation), _lib.ptr(opacity),
                _lib.ptr(g_means), _lib.ptr(g_opac), _lib.ptr(g_scales), _lib.ptr(g_rots), _lib.ptr(acc_xyz),
                _lib.ptr(g_T), _lib.ptr(acc_opacity), _lib.ptr(acc_scaling), _lib.ptr(acc_rotation),
                C.c_void_p(stream)))
            if not six:
                g_T = g_means if need[1] else None
            if ON_GAUSSIAN_GRADS is not None:
                ON_GAUSSIAN_GRADS()
            return (None, g_T, None, g_scales if need[3] else None, None, g_rots if need[5] else None, None, None,
                    None)
        g_xyz = new(N, 3) if (six and need[0]) else None
        g_T = new(N, 4, 4) if (six and need[1]) else None
        g_scaling = new(N, 3) if need[2] else None
        g_rotation = new(N, 4) if need[4] else None
        g_opacity = new(N, 1) if need[6] else None
        stream = torch.cuda.current_stream().cuda_stream
        _lib.check(lib.d3g_gaussians_prepare_backward(
            N, int(six), _lib.ptr(xyz), _lib.ptr(d_xyz), _lib.ptr(scaling), _lib.ptr(rotation), _lib.ptr(opacity),
            _lib.ptr(g_means), _lib.ptr(g_opac), _lib.ptr(g_scales), _lib.ptr(g_rots), _lib.ptr(g_xyz),
            _lib.ptr(g_T), _lib.ptr(g_opacity), _lib.ptr(g_scaling), _lib.ptr(g_rotation), C.c_void_p(stream)))
        if not six:  # additive: both operands receive the incoming cotangent itself
            g_xyz = g_means if need[0] else None
            g_T = g_means if need[1] else None
        return (g_xyz, g_T, g_scaling, g_scales if need[3] else None, g_rotation, g_rots if need[5] else None,
                g_opacity, None, None)


def _can_fuse_prepare(pc, d_xyz, d_rotation, d_scaling, is_6dof):
    if not FUSED_PREPARE:
        return False
    if not (torch.is_tensor(d_xyz) and torch.is_tensor(d_rotation) and torch.is_tensor(d_scaling)):
        return False
    if (getattr(pc, "scaling_activation", None) is not torch.exp
            or getattr(pc, "opacity_activation", None) is not torch.sigmoid
            or getattr(pc, "rotation_activation", None) is not torch.nn.functional.normalize):
        return False
    raw = [getattr(pc, n, None) for n in ("_xyz", "_scaling", "_rotation", "_opacity")]
    if any(not torch.is_tensor(t) for t in raw):
        return False
    N = raw[0].shape[0]
    shapes = [(N, 3), (N, 3), (N, 4), (N, 1)]
    want_d = [(N, 4, 4) if is_6dof else (N, 3), (N, 4), (N, 3)]
    tensors = raw + [d_xyz, d_rotation, d_scaling]
    for t, shp in zip(tensors, shapes + want_d):
        if tuple(t.shape) != shp or t.dtype != torch.float32 or not t.is_cuda or not t.is_contiguous():
            return False
    return True


def _grad_sinks(pc):
    """The six .grad tensors FUSED_GRAD_ACCUMULATION adds into — (xyz, scaling, rotation, opacity, features_dc,
    features_rest) — or None when `pc` does not qualify (then render() takes the autograd path)."""
    if not (FUSED_GRAD_ACCUMULATION and torch.is_grad_enabled()):
        return None
    names = ("_xyz", "_scaling", "_rotation", "_opacity", "_features_dc", "_features_rest")
    params = [getattr(pc, n, None) for n in names]
    if not torch.is_tensor(params[0]):
        return None
    N = params[0].shape[0]
    shapes = [(N, 3), (N, 3), (N, 4), (N, 1), (N, 1, 3), (N, 15, 3)]
    for p, shp in zip(params, shapes):
        if not torch.is_tensor(p) or not p.requires_grad or not p.is_leaf or tuple(p.shape) != shp:
            return None
        g = p.grad
        if (g is None or g.dtype != torch.float32 or not g.is_contiguous() or g.device != p.device
                or p.dtype != torch.float32 or not p.is_contiguous() or not p.is_cuda):
            return None
    rest = params[5]
    if rest.data_ptr() % 16 or rest.grad.data_ptr() % 16 or getattr(pc, "max_sh_degree", 3) != 3:
        return None
    return tuple(p.grad for p in params)


def eval_sh(deg, sh, dirs):
    """SH polynomial of utils/sh_utils.py:57-112 (used only by the convert_SHs_python pipe flag)."""
    C0, C1 = 0.28209479177387814, 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
          1.445305721320277, -0.5900435899266435]
    result = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        result = result - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            result = (result + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5]
                      + C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + C2[3] * xz * sh[..., 7]
                      + C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                result = (result + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10]
                          + C3[2] * y * (4 * zz - xx - yy) * sh[..., 11]
                          + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12]
                          + C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14]
                          + C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return result


def _to_homogenous(v):
    return torch.cat([v, torch.ones_like(v[..., :1])], dim=-1)


def _from_homogenous(v):
    return v[..., :3] / v[..., -1:]


def _build_covariance(scaling, modifier, rotation):
    """scene/gaussian_model.py:28-32 (+ utils/general_utils.py:130-163), device-agnostic."""
    q = rotation / rotation.norm(dim=1, keepdim=True)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=-1).reshape(-1, 3, 3)
    L = R * (modifier * scaling).unsqueeze(1)
    S = L @ L.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], dim=-1)


def render(viewpoint_camera, pc, pipe, bg_color, d_xyz, d_rotation, d_scaling, is_6dof=False,
           scaling_modifier=1.0, override_color=None, direct_compute=False):
    xyz = pc.get_xyz
    # (the reference builds these as `zeros_like(...) + 0` and retains the gradient of the sum; leaf tensors
    # behave identically for every consumer — .grad is populated by backward — without the two add launches)
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device)
    screenspace_points_densify = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device)

    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height),
        image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx,
        tanfovy=tanfovy,
        bg=bg_color,
        scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center,
        prefiltered=False,
        debug=getattr(pipe, "debug", False),
    )
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    fused = (not direct_compute and not getattr(pipe, "compute_cov3D_python", False)
             and _can_fuse_prepare(pc, d_xyz, d_rotation, d_scaling, is_6dof))
    # (the SH-coefficient sinks need the rasterizer's SH path: not with the convert_SHs_python pipe flag)
    sinks = _grad_sinks(pc) if (fused and not getattr(pipe, "convert_SHs_python", False)) else None
    if fused:
        means3D, opacity, scales, rotations = _PrepareGaussians.apply(
            pc._xyz, d_xyz, pc._scaling, d_scaling, pc._rotation, d_rotation, pc._opacity, bool(is_6dof),
            None if sinks is None else sinks[:4])
    elif direct_compute:
        means3D = d_xyz
    elif is_6dof:
        if torch.is_tensor(d_xyz) is False:
            means3D = xyz
        else:
            means3D = _from_homogenous(torch.bmm(d_xyz, _to_homogenous(xyz).unsqueeze(-1)).squeeze(-1))
    else:
        means3D = xyz + d_xyz
    cov3D_precomp = None
    if not fused:
        opacity = pc.get_opacity
        scales = rotations = None
    if fused:
        pass  # scales / rotations / opacity came out of the fused kernel
    elif getattr(pipe, "compute_cov3D_python", False):
        if hasattr(pc, "get_covariance"):
            cov3D_precomp = pc.get_covariance(scaling_modifier)
        else:
            cov3D_precomp = _build_covariance(pc.get_scaling, scaling_modifier, pc._rotation)
    elif d_scaling is None and d_rotation is None:
        scales, rotations = pc.get_scaling, pc.get_rotation
    elif direct_compute:
        scales, rotations = d_scaling, d_rotation
    else:
        scales = pc.get_scaling + d_scaling
        rotations = pc.get_rotation + d_rotation

    shs = colors_precomp = None
    if sinks is not None:
        pass  # the rasterizer reads _features_dc / _features_rest in place: no `get_features` concatenation
    elif colors_precomp is None:  # always true, as in the reference (:103)
        if getattr(pipe, "convert_SHs_python", False):
            feats = pc.get_features
            shs_view = feats.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            dir_pp = xyz - viewpoint_camera.camera_center.repeat(feats.shape[0], 1)
            dir_pp_normalized = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            sh2rgb = eval_sh(pc.active_sh_degree, shs_view, dir_pp_normalized)
            colors_precomp = torch.clamp_min(sh2rgb + 0.5, 0.0)
        else:
            shs = pc.get_features
    else:
        colors_precomp = override_color

    if sinks is not None:
        # same rasterizer, SH coefficients read from (and their gradients added to) the two parameters in place
        rendered_image, radii, depth = rasterize_gaussians(
            means3D, screenspace_points, screenspace_points_densify, None, None, opacity, scales, rotations, None,
            raster_settings, (pc._features_dc, pc._features_rest, sinks[4], sinks[5]))
    else:
        rendered_image, radii, depth = rasterizer(
            means3D=means3D, means2D=screenspace_points, means2D_densify=screenspace_points_densify, shs=shs,
            colors_precomp=colors_precomp, opacities=opacity, scales=scales, rotations=rotations,
            cov3D_precomp=cov3D_precomp)

    return {"render": rendered_image,
            "viewspace_points": screenspace_points,
            "viewspace_points_densify": screenspace_points_densify,
            "visibility_filter": radii > 0,
            "radii": radii,
            "depth": depth}
