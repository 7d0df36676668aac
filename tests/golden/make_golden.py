"""Generates the golden fixtures in this directory by IMPORTING THE REFERENCE'S OWN PYTHON CODE.

Runs only in the build container (needs /root/reference); the fixtures it writes are committed and are
what the CPU test-suite and the GPU parity tests compare against.  Nothing under tests/, bench.py or
smoke() reads /root/reference at run time.

    python tests/golden/make_golden.py

Fixtures
  mlp_{blender,real,6dof}.npz  reference DeformNetworkBaseline / DeformNetwork (utils/time_utils.py)
                               forward outputs + parameter-gradient digests for seeded inputs
  sh_eval.npz                  reference eval_sh (utils/sh_utils.py:57-112) colours
  camera.npz                   reference getWorld2View2 / getProjectionMatrix / pose_spherical matrices
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import mlp_ref  # noqa: E402  (only for the seeded parameter generator)

from utils.time_utils import DeformNetwork, DeformNetworkBaseline  # noqa: E402  reference
from utils.sh_utils import eval_sh  # noqa: E402  reference
from utils.graphics_utils import getWorld2View2, getProjectionMatrix  # noqa: E402  reference
from utils.pose_utils import pose_spherical  # noqa: E402  reference


def digest(g, probe_seed):
    """Small, order-sensitive summary of a gradient tensor: norm, projection on a seeded probe, head."""
    g = np.asarray(g, np.float64)
    probe = np.random.default_rng(probe_seed).standard_normal(g.shape)
    return np.array([np.sqrt((g * g).sum()), (g * probe).sum(), np.abs(g).max()]), g.reshape(-1)[:48].copy()


def mlp_case(name, is_blender, is_6dof, baseline, seed, N=257, uniform_t=True):
    torch.manual_seed(0)
    cls = DeformNetworkBaseline if baseline else DeformNetwork
    net = cls(is_blender=is_blender, is_6dof=is_6dof).double()
    params = mlp_ref.seeded_params(seed, is_blender, is_6dof, head_scale=0.05 if is_6dof else 1.0)
    sd = net.state_dict()
    assert set(sd.keys()) == set(params.keys()), (sorted(sd.keys()), sorted(params.keys()))
    for k in sd:
        assert tuple(sd[k].shape) == params[k].shape, k
    net.load_state_dict({k: torch.tensor(v, dtype=torch.float64) for k, v in params.items()})
    rng = np.random.default_rng(seed + 100)
    x = rng.uniform(-1.3, 1.3, size=(N, 3)).astype(np.float32)
    if uniform_t:
        t = np.full((N, 1), 0.37, np.float32)
    else:
        t = rng.uniform(0, 1, size=(N, 1)).astype(np.float32)
    xt = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    tt = torch.tensor(t, dtype=torch.float64, requires_grad=True)
    d_xyz, rot, sc = net(xt, tt)
    g_dxyz = rng.standard_normal(tuple(d_xyz.shape)).astype(np.float32)
    loss = (d_xyz * torch.tensor(g_dxyz, dtype=torch.float64)).sum()
    out = dict(x=x, t=t, seed=np.int64(seed), d_xyz=d_xyz.detach().numpy(), g_dxyz=g_dxyz)
    if baseline:
        g_rot = rng.standard_normal((N, 4)).astype(np.float32)
        g_sc = rng.standard_normal((N, 3)).astype(np.float32)
        loss = loss + (rot * torch.tensor(g_rot, dtype=torch.float64)).sum() + (sc * torch.tensor(g_sc, dtype=torch.float64)).sum()
        out.update(rot=rot.detach().numpy(), scale=sc.detach().numpy(), g_rot=g_rot, g_scale=g_sc)
    loss.backward()
    for i, (k, p) in enumerate(sorted(net.named_parameters())):
        if p.grad is None:
            continue
        d, head = digest(p.grad.numpy(), 1000 + i)
        out["grad_digest/" + k] = d
        out["grad_head/" + k] = head
        if p.grad.numel() <= 4096:
            out["grad_full/" + k] = p.grad.numpy().copy()
    out["grad_x"] = xt.grad.numpy().copy()
    out["grad_t"] = tt.grad.numpy().copy()
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "d_xyz", out["d_xyz"].shape, "max|d_xyz|", np.abs(out["d_xyz"]).max())


def sh_case():
    rng = np.random.default_rng(7)
    N = 96
    sh = rng.standard_normal((N, 16, 3)).astype(np.float32) * 0.4
    dirs = rng.standard_normal((N, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    out = {"sh": sh, "dirs": dirs}
    for deg in range(4):
        c = eval_sh(deg, torch.tensor(sh).transpose(1, 2).double(), torch.tensor(dirs).double())
        out["rgb_deg%d" % deg] = torch.clamp_min(c + 0.5, 0.0).numpy()
    np.savez_compressed(os.path.join(HERE, "sh_eval.npz"), **out)
    print("sh_eval ok")


def camera_case():
    out = {}
    for i, (theta, W, H) in enumerate([(0.0, 800, 800), (37.0, 400, 400), (-120.0, 1920, 1080)]):
        pose = pose_spherical(theta, -30.0, 4.0)
        matrix = np.linalg.inv(np.array(pose))
        R = -np.transpose(matrix[:3, :3])
        R[:, 0] = -R[:, 0]
        T = -matrix[:3, 3]
        fovx = 0.6911112
        fovy = fovx if W == H else 2 * np.arctan(H / (2 * (W / (2 * np.tan(fovx / 2)))))
        wvt = torch.tensor(getWorld2View2(R, T)).transpose(0, 1)
        proj = getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fovx, fovY=fovy).transpose(0, 1)
        full = wvt.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0)
        out["args%d" % i] = np.array([theta, W, H, fovx, fovy])
        out["wvt%d" % i] = wvt.numpy()
        out["full%d" % i] = full.numpy()
        out["center%d" % i] = wvt.inverse()[3, :3].numpy()
    np.savez_compressed(os.path.join(HERE, "camera.npz"), **out)
    print("camera ok")


if __name__ == "__main__":
    mlp_case("mlp_blender.npz", True, False, True, seed=11)
    mlp_case("mlp_blender_dn.npz", True, False, False, seed=12)
    mlp_case("mlp_real.npz", False, False, True, seed=13, uniform_t=False)
    mlp_case("mlp_6dof.npz", True, True, True, seed=14)
    mlp_case("mlp_blender_rowt.npz", True, False, True, seed=15, uniform_t=False)
    sh_case()
    camera_case()
