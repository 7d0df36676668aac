"""Synthetic scenes and cameras for parity tests and bench.py.

Everything is generated on the CPU from a seeded torch.Generator so that the build container, the
GPU box and every rank see bit-identical inputs; callers move the tensors to the device.

Camera maths follows the reference: getWorld2View2 / getProjectionMatrix
(utils/graphics_utils.py:42-76), Camera.__init__ (scene/cameras.py:49-61: znear=0.01, zfar=100,
matrices stored transposed), pose_spherical (utils/pose_utils.py:59-64) with the render.py:275-278
c2w -> (R, T) conversion.  Gaussian activations follow scene/gaussian_model.py:58-78.
"""
import math
from types import SimpleNamespace

import numpy as np
import torch


def _world2view(R, t):
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    return np.float32(Rt)


def _projection(znear, zfar, fovX, fovY):
    tanY, tanX = math.tan(fovY / 2), math.tan(fovX / 2)
    top, right = tanY * znear, tanX * znear
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def _pose_spherical(theta, phi, radius):
    tr = np.eye(4, dtype=np.float32)
    tr[2, 3] = radius
    ph = phi / 180.0 * np.pi
    rp = np.array([[1, 0, 0, 0], [0, np.cos(ph), -np.sin(ph), 0], [0, np.sin(ph), np.cos(ph), 0],
                   [0, 0, 0, 1]], dtype=np.float32)
    th = theta / 180.0 * np.pi
    rt = np.array([[np.cos(th), 0, -np.sin(th), 0], [0, 1, 0, 0], [np.sin(th), 0, np.cos(th), 0],
                   [0, 0, 0, 1]], dtype=np.float32)
    c2w = rt @ rp @ tr
    flip = np.array([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]], dtype=np.float32)
    return flip @ c2w


def make_camera(width, height, theta_deg=0.0, phi_deg=-30.0, radius=4.0, fovx=0.6911112, fovy=None,
                fid=0.0):
    """A MiniCam-like namespace with the attributes render() reads (scene/cameras.py:78-89)."""
    if fovy is None:
        if width == height:
            fovy = fovx
        else:  # focal2fov(fov2focal(fovx, W), H), utils/graphics_utils.py:79-84
            focal = width / (2 * math.tan(fovx / 2))
            fovy = 2 * math.atan(height / (2 * focal))
    c2w = _pose_spherical(theta_deg, phi_deg, radius).astype(np.float64)
    mat = np.linalg.inv(c2w)
    R = -np.transpose(mat[:3, :3])
    R[:, 0] = -R[:, 0]
    T = -mat[:3, 3]
    wvt = torch.tensor(_world2view(R, T)).transpose(0, 1).contiguous()
    projm = _projection(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = wvt.unsqueeze(0).bmm(projm.unsqueeze(0)).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    return SimpleNamespace(image_width=int(width), image_height=int(height), FoVx=float(fovx),
                           FoVy=float(fovy), znear=0.01, zfar=100.0, world_view_transform=wvt,
                           full_proj_transform=full, camera_center=center,
                           fid=torch.tensor([float(fid)]))


def camera_to(cam, device):
    c = SimpleNamespace(**vars(cam))
    for k in ("world_view_transform", "full_proj_transform", "camera_center", "fid"):
        setattr(c, k, getattr(cam, k).to(device))
    return c


class SynthGaussians:
    """Minimal stand-in for scene/gaussian_model.py::GaussianModel: the six raw parameter tensors and
    the activated getters render() consumes (gaussian_model.py:58-78)."""

    def __init__(self, N, seed=0, sh_degree=3, kind="uniform", device="cpu"):
        g = torch.Generator().manual_seed(seed)
        if kind == "blob":  # config C1, "hook"-shaped object-centric blob
            xyz = (0.45 * torch.randn(N, 3, generator=g)).clamp(-1.3, 1.3)
        else:  # the reference's own random init, scene/dataset_readers.py:286-290
            xyz = torch.rand(N, 3, generator=g) * 2.6 - 1.3
        mu = math.log(0.012 * (1e5 / N) ** (1.0 / 3.0))
        self._xyz = xyz.to(device)
        self._scaling = (mu + 0.5 * torch.randn(N, 3, generator=g)).to(device)
        self._rotation = torch.randn(N, 4, generator=g).to(device)
        self._opacity = (2.0 * torch.randn(N, 1, generator=g)).to(device)
        M = (sh_degree + 1) ** 2
        self._features_dc = (0.5 * torch.randn(N, 1, 3, generator=g)).to(device)
        self._features_rest = (0.1 * torch.randn(N, M - 1, 3, generator=g)).to(device)
        self.active_sh_degree = sh_degree
        self.max_sh_degree = sh_degree
        # GaussianModel.setup_functions, scene/gaussian_model.py:58-78
        self.scaling_activation = torch.exp
        self.opacity_activation = torch.sigmoid
        self.rotation_activation = torch.nn.functional.normalize

    def parameters(self):
        return [self._xyz, self._features_dc, self._features_rest, self._opacity, self._scaling,
                self._rotation]

    def requires_grad_(self, flag=True):
        for p in self.parameters():
            p.requires_grad_(flag)
        return self

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_scaling(self):
        return self.scaling_activation(self._scaling)

    @property
    def get_rotation(self):
        return self.rotation_activation(self._rotation)

    @property
    def get_opacity(self):
        return self.opacity_activation(self._opacity)

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)


def image_cotangent(H, W, seed=1):
    """Kernel-level cotangent dL/dimage = randn(3,H,W)/(3HW)."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(3, H, W, generator=g) / (3.0 * H * W)
