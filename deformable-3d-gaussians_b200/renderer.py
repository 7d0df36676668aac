"""Drop-in for `gaussian_renderer.render` (gaussian_renderer/__init__.py:32-133).

Same signature, same branching (including the quirks listed in SURVEY.md §7: un-normalised
`get_rotation + d_rotation`, deltas used verbatim under `direct_compute`, unreachable `override_color`)
and the same returned dict.  The rasterizer behind it is this package's C-ABI one.
"""
import math

import torch

from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer


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
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    screenspace_points_densify = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=xyz.device) + 0
    try:
        screenspace_points.retain_grad()
        screenspace_points_densify.retain_grad()
    except Exception:
        pass

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

    if direct_compute:
        means3D = d_xyz
    elif is_6dof:
        if torch.is_tensor(d_xyz) is False:
            means3D = xyz
        else:
            means3D = _from_homogenous(torch.bmm(d_xyz, _to_homogenous(xyz).unsqueeze(-1)).squeeze(-1))
    else:
        means3D = xyz + d_xyz
    opacity = pc.get_opacity

    scales = rotations = cov3D_precomp = None
    if getattr(pipe, "compute_cov3D_python", False):
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
    if colors_precomp is None:  # always true, as in the reference (:103)
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
