"""Pins oracle/mlp_ref.py (numpy restatement) to golden vectors produced by the REFERENCE's own torch
module (tests/golden/make_golden.py imports /root/reference/utils/time_utils.py)."""
import os

import numpy as np
import pytest

from oracle import mlp_ref

GOLD = os.path.join(os.path.dirname(__file__), "golden")

CASES = [("mlp_blender.npz", True, False, True, 1.0), ("mlp_blender_dn.npz", True, False, False, 1.0),
         ("mlp_real.npz", False, False, True, 1.0), ("mlp_6dof.npz", True, True, True, 0.05),
         ("mlp_blender_rowt.npz", True, False, True, 1.0)]


def _digest(g, probe_seed):
    g = np.asarray(g, np.float64)
    probe = np.random.default_rng(probe_seed).standard_normal(g.shape)
    return np.array([np.sqrt((g * g).sum()), (g * probe).sum(), np.abs(g).max()])


@pytest.mark.parametrize("fname,is_blender,is_6dof,baseline,head_scale", CASES)
def test_oracle_matches_reference_golden(fname, is_blender, is_6dof, baseline, head_scale):
    z = np.load(os.path.join(GOLD, fname))
    params = mlp_ref.seeded_params(int(z["seed"]), is_blender, is_6dof, head_scale=head_scale)
    net = mlp_ref.DeformRef(params, is_blender, is_6dof, baseline, dtype=np.float64)
    d_xyz, rot, sc = net.forward(z["x"], z["t"])
    np.testing.assert_allclose(d_xyz, z["d_xyz"], rtol=1e-9, atol=1e-11)
    if baseline:
        np.testing.assert_allclose(rot, z["rot"], rtol=1e-9, atol=1e-11)
        np.testing.assert_allclose(sc, z["scale"], rtol=1e-9, atol=1e-11)
    grads = net.backward(z["g_dxyz"], z["g_rot"] if baseline else None, z["g_scale"] if baseline else None,
                         want_input_grads=True)
    names = sorted(k[len("grad_digest/"):] for k in z.files if k.startswith("grad_digest/"))
    assert names, "fixture has no gradients"
    all_names = sorted(mlp_ref.param_shapes(is_blender, is_6dof).keys())
    for k in names:
        i = all_names.index(k)
        want = z["grad_digest/" + k]
        got = _digest(grads[k], 1000 + i)
        np.testing.assert_allclose(got, want, rtol=1e-8, atol=1e-10, err_msg=k)
        np.testing.assert_allclose(grads[k].reshape(-1)[:48], z["grad_head/" + k], rtol=1e-8, atol=1e-10, err_msg=k)
        if "grad_full/" + k in z.files:
            np.testing.assert_allclose(grads[k], z["grad_full/" + k], rtol=1e-8, atol=1e-10, err_msg=k)
    np.testing.assert_allclose(grads["_x"], z["grad_x"], rtol=1e-8, atol=1e-9)
    np.testing.assert_allclose(grads["_t"], z["grad_t"], rtol=1e-8, atol=1e-9)


def test_oracle_fp32_close_to_fp64():
    z = np.load(os.path.join(GOLD, "mlp_blender.npz"))
    params = mlp_ref.seeded_params(int(z["seed"]))
    d32 = mlp_ref.DeformRef(params, dtype=np.float32).forward(z["x"], z["t"])[0]
    assert np.abs(d32 - z["d_xyz"]).max() / np.abs(z["d_xyz"]).max() < 1e-5


def test_param_shapes_count():
    # 522,280 parameters for is_blender (verified against the reference module)
    n = sum(int(np.prod(s)) for s in mlp_ref.param_shapes(True, False).values())
    assert n == 522280
