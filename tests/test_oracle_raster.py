"""CPU tests of the rasterizer oracle (oracle/raster_ref.c): analytic gradients vs fp64 finite
differences, structural invariants of the binning stage, SH colours vs the reference's eval_sh
(golden), hand-computed single-Gaussian case."""
import math
import os

import numpy as np
import pytest

from conftest import pkg
from oracle import raster_oracle as ro

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def make_scene(N, H, W, seed=0, dtype=np.float32, theta=20.0, scale_mul=1.0, kind="uniform", bg=(0.1, 0.2, 0.3)):
    synth = pkg("synth")
    g = synth.SynthGaussians(N, seed=seed, kind=kind)
    cam = synth.make_camera(W, H, theta_deg=theta)
    cfg = ro.make_cfg(H, W, math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5), bg, 1.0,
                      cam.world_view_transform.numpy(), cam.full_proj_transform.numpy(),
                      cam.camera_center.numpy(), 3, 16, 2, dtype)
    inp = dict(means3D=g.get_xyz.numpy(), shs=g.get_features.numpy(), opacities=g.get_opacity.numpy(),
               scales=g.get_scaling.numpy() * scale_mul, rots=g.get_rotation.numpy())
    return cfg, inp, cam


def run_fwd(cfg, inp, dtype=np.float32):
    return ro.forward(cfg, inp["means3D"], inp["shs"], None, inp["opacities"], inp["scales"], inp["rots"], None,
                      dtype=dtype)


def test_binning_invariants():
    cfg, inp, _ = make_scene(3000, 160, 208)
    f = run_fwd(cfg, inp)
    R = f["R"]
    assert R == int(f["tiles_touched"].astype(np.int64).sum())
    keys = f["keys"][:R]
    assert np.all(keys[1:] >= keys[:-1]), "sorted keys must be monotone"
    # the sort is a permutation of the emitted pairs
    assert np.array_equal(np.sort(f["keys_unsorted"][:R]), keys)
    # ranges partition [0,R) over non-empty tiles
    rg = f["ranges"].astype(np.int64)
    ne = rg[rg[:, 1] > rg[:, 0]]
    assert ne[:, 0].min() == 0 and ne[:, 1].max() == R
    assert np.array_equal(np.sort(ne[:, 0])[1:], np.sort(ne[:, 1])[:-1])
    tiles = (keys >> np.uint64(32)).astype(np.int64)
    for t in np.unique(tiles)[:50]:
        assert np.all(tiles[rg[t, 0]:rg[t, 1]] == t)
    # stability: equal keys keep ascending Gaussian index
    same = keys[1:] == keys[:-1]
    pl = f["point_list"][:R].astype(np.int64)
    assert np.all(pl[1:][same] > pl[:-1][same])
    # radii > 0 <=> tiles_touched > 0
    assert np.array_equal(f["radii"] > 0, f["tiles_touched"] > 0)


def test_single_gaussian_known_answer():
    """One isotropic Gaussian on the optical axis: radius, tile rect and centre pixel by hand."""
    synth = pkg("synth")
    H = W = 64
    cam = synth.make_camera(W, H, theta_deg=0.0, phi_deg=0.0, radius=4.0)
    tanf = math.tan(cam.FoVx * 0.5)
    cfg = ro.make_cfg(H, W, tanf, tanf, (0, 0, 0), 1.0, cam.world_view_transform.numpy(),
                      cam.full_proj_transform.numpy(), cam.camera_center.numpy(), 0, 1, 2, np.float64)
    s = 0.05
    f = ro.forward(cfg, np.zeros((1, 3)), np.full((1, 1, 3), 1.0), None, np.array([[0.9]]), np.full((1, 3), s),
                   np.array([[1.0, 0, 0, 0]]), None, dtype=np.float64)
    focal = W / (2 * tanf)
    var = (focal * s / 4.0) ** 2 + 0.3
    # isotropic: mid^2 - det = 0, so the eigenvalue floor max(0.1, .) applies
    assert f["radii"][0] == math.ceil(3 * math.sqrt(var + math.sqrt(0.1)))
    np.testing.assert_allclose(f["means2D"][0], [(W - 1) / 2, (H - 1) / 2], atol=1e-9)
    np.testing.assert_allclose(f["depths"][0], 4.0, atol=1e-6)
    np.testing.assert_allclose(f["conic_opacity"][0], [1 / var, 0, 1 / var, 0.9], atol=1e-9)
    r = f["radii"][0]
    lo, hi = int((31.5 - r) / 16), int((31.5 + r + 15) / 16)
    assert f["tiles_touched"][0] == (hi - lo) ** 2
    # centre pixel (31,31): offset 0.5 px in x and y
    a = min(0.99, 0.9 * math.exp(-0.5 * (0.5 / var)))
    col = (0.28209479177387814 * 1.0 + 0.5) * a
    np.testing.assert_allclose(f["color"][:, 31, 31], [col] * 3, rtol=1e-9)
    np.testing.assert_allclose(f["depth"][0, 31, 31], 4.0 * a, rtol=1e-6)
    np.testing.assert_allclose(f["final_T"][31, 31], 1 - a, rtol=1e-9)


def test_fp32_close_to_fp64_image():
    cfg, inp, _ = make_scene(1500, 96, 96, seed=2, scale_mul=2.0)
    cfg64, _, _ = make_scene(1500, 96, 96, seed=2, scale_mul=2.0, dtype=np.float64)
    f32 = run_fwd(cfg, inp)
    f64 = run_fwd(cfg64, inp, np.float64)
    # blending decisions can flip at thresholds between precisions on isolated pixels; the bulk agrees
    diff = np.abs(f32["color"] - f64["color"])
    assert np.quantile(diff, 0.999) < 5e-5
    assert np.mean(f32["radii"] == f64["radii"]) > 0.999


@pytest.mark.parametrize("with_depth", [False, True])
def test_gradients_vs_finite_differences(with_depth):
    cfg, inp, _ = make_scene(250, 64, 64, seed=3, dtype=np.float64, scale_mul=3.0)
    rng = np.random.default_rng(0)
    dimg = rng.standard_normal((3, 64, 64))
    ddep = rng.standard_normal((1, 64, 64)) * 0.1 if with_depth else None

    def loss(i):
        f = run_fwd(cfg, i, np.float64)
        return f, float((f["color"] * dimg).sum() + (0 if ddep is None else (f["depth"] * ddep).sum()))

    f0, _ = loss(inp)
    g = ro.backward(cfg, f0, dimg, ddep, dtype=np.float64)
    names = dict(means3D="dL_dmeans3D", shs="dL_dsh", opacities="dL_dopacity", scales="dL_dscales", rots="dL_drots")
    vis = np.where(f0["radii"] > 0)[0]
    assert len(vis) > 100
    for k, gk in names.items():
        scale = np.abs(g[gk]).max()
        for _ in range(12):
            i = vis[rng.integers(len(vis))]
            x = inp[k].astype(np.float64)
            idx = (i,) + tuple(rng.integers(s) for s in x.shape[1:])
            eps = 1e-6 * max(1.0, abs(x[idx]))
            xp, xm = x.copy(), x.copy()
            xp[idx] += eps
            xm[idx] -= eps
            fd = (loss({**inp, k: xp})[1] - loss({**inp, k: xm})[1]) / (2 * eps)
            assert abs(fd - g[gk][idx]) <= 1e-5 * scale + 1e-7, (k, idx, fd, g[gk][idx])


def test_densify_modes():
    cfg, inp, _ = make_scene(400, 64, 64, seed=4, dtype=np.float64, scale_mul=3.0)
    f = run_fwd(cfg, inp, np.float64)
    dimg = np.random.default_rng(1).standard_normal((3, 64, 64))
    cfg.densify_mode = 1
    g1 = ro.backward(cfg, f, dimg, None, np.float64)
    np.testing.assert_allclose(g1["dL_dmean2D_densify"], g1["dL_dmean2D"])
    cfg.densify_mode = 2
    g2 = ro.backward(cfg, f, dimg, None, np.float64)
    assert np.all(g2["dL_dmean2D_densify"] >= np.abs(g2["dL_dmean2D"]) - 1e-12)
    cfg.densify_mode = 3
    g3 = ro.backward(cfg, f, dimg, None, np.float64)
    assert np.all(g3["dL_dmean2D_densify"][:, 0] >= np.linalg.norm(g3["dL_dmean2D"], axis=1) - 1e-12)
    assert np.all(g3["dL_dmean2D_densify"][:, 1] == 0)


def test_sh_colour_matches_reference_eval_sh():
    """Camera at the origin looking down +z with a very wide frustum, so that a point placed at 3*dir
    is visible whenever dir.z > 0.1 and its SH direction is exactly the fixture's `dir`."""
    z = np.load(os.path.join(GOLD, "sh_eval.npz"))
    dirs = z["dirs"].astype(np.float64)
    xyz = 3.0 * dirs
    N = xyz.shape[0]
    view = np.eye(4)
    proj = np.zeros(16)
    proj[0], proj[5], proj[10], proj[11] = 0.2, 0.2, 1.0, 1.0  # flat m[4*col+row]: w = z
    for deg in range(4):
        cfg = ro.make_cfg(512, 512, 5.0, 5.0, (0, 0, 0), 1.0, view, proj, (0, 0, 0), deg, 16, 2, np.float64)
        f = ro.forward(cfg, xyz, z["sh"], None, np.full((N, 1), 0.5), np.full((N, 3), 0.05),
                       np.tile([1.0, 0, 0, 0], (N, 1)), None, dtype=np.float64)
        vis = f["radii"] > 0
        assert vis.sum() > 20
        np.testing.assert_allclose(f["rgb"][vis], z["rgb_deg%d" % deg][vis], atol=2e-6)
        # fp32 spec path agrees too
        cfg32 = ro.make_cfg(512, 512, 5.0, 5.0, (0, 0, 0), 1.0, view, proj, (0, 0, 0), deg, 16, 2, np.float32)
        f32 = ro.forward(cfg32, xyz, z["sh"], None, np.full((N, 1), 0.5), np.full((N, 3), 0.05),
                         np.tile([1.0, 0, 0, 0], (N, 1)), None, dtype=np.float32)
        v32 = f32["radii"] > 0
        np.testing.assert_allclose(f32["rgb"][v32], z["rgb_deg%d" % deg][v32], atol=5e-6)


def test_camera_matches_reference():
    z = np.load(os.path.join(GOLD, "camera.npz"))
    synth = pkg("synth")
    for i in range(3):
        theta, W, H, fovx, fovy = z["args%d" % i]
        cam = synth.make_camera(int(W), int(H), theta_deg=float(theta))
        np.testing.assert_allclose(cam.world_view_transform.numpy(), z["wvt%d" % i], atol=1e-6)
        np.testing.assert_allclose(cam.full_proj_transform.numpy(), z["full%d" % i], atol=1e-5)
        np.testing.assert_allclose(cam.camera_center.numpy(), z["center%d" % i], atol=1e-5)
        assert abs(cam.FoVy - fovy) < 1e-9


def test_empty_scene_and_offscreen():
    cfg, inp, _ = make_scene(50, 32, 48, seed=5)
    far = {k: v.copy() for k, v in inp.items()}
    far["means3D"][:] += 100.0  # everything behind / outside
    f = run_fwd(cfg, far)
    assert f["R"] == 0 and np.all(f["radii"] == 0)
    np.testing.assert_allclose(f["color"][0], 0.1, atol=1e-7)
    np.testing.assert_allclose(f["final_T"], 1.0)
    g = ro.backward(cfg, f, np.ones((3, 32, 48), np.float32))
    assert all(np.all(v == 0) for k, v in g.items() if v is not None)


def test_det_exp_accuracy():
    """The pinned deterministic exp of the fp32 spec vs true exp on the range the blend can reach
    (alpha >= 1/255 with opacity <= 1 means power >= ln(1/255) = -5.54)."""
    import ctypes as C
    L = ro.lib()
    x = np.linspace(-6.0, 0.0, 200001).astype(np.float32)
    y = np.zeros_like(x)
    L.det_exp_array_f32(x.size, x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p))
    ref = np.exp(x.astype(np.float64))
    assert np.max(np.abs(y - ref) / ref) < 3e-7
    # far tail: monotone to zero, never negative / NaN
    xt = np.array([-20.0, -80.0, -87.5, -1000.0], np.float32)
    yt = np.zeros_like(xt)
    L.det_exp_array_f32(xt.size, xt.ctypes.data_as(C.c_void_p), yt.ctypes.data_as(C.c_void_p))
    assert np.all(yt >= 0) and yt[-1] == 0 and abs(yt[0] - np.exp(-20.0)) / np.exp(-20.0) < 1e-6
