"""GPU parity of the fused deformation MLP (through the C ABI) against the numpy oracle, which is itself
pinned to the reference's torch module (tests/test_oracle_mlp.py).  Tolerance from BASELINE.json:
deformation offsets and gradients within 1e-3 relative, measured per tensor as max|a-b| / max|b|
(TF32 tensor-core inputs, fp32 accumulation)."""
import os

import numpy as np
import pytest
import torch

from conftest import pkg
from oracle import mlp_ref

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-3


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def _make_net(is_blender, is_6dof, baseline, params):
    deform = pkg("deform")
    cls = deform.DeformNetworkBaseline if baseline else deform.DeformNetwork
    net = cls(is_blender=is_blender, is_6dof=is_6dof)
    sd = net.state_dict()
    assert set(sd.keys()) == set(params.keys())
    net.load_state_dict({k: torch.tensor(v) for k, v in params.items()})
    net.debug = True
    return net.cuda()


CASES = [("mlp_blender.npz", True, False, True, 1.0), ("mlp_blender_dn.npz", True, False, False, 1.0),
         ("mlp_real.npz", False, False, True, 1.0), ("mlp_6dof.npz", True, True, True, 0.05),
         ("mlp_blender_rowt.npz", True, False, True, 1.0)]


@pytest.mark.parametrize("fname,is_blender,is_6dof,baseline,head_scale", CASES)
def test_forward_matches_reference_golden(fname, is_blender, is_6dof, baseline, head_scale):
    z = np.load(os.path.join(GOLD, fname))
    params = mlp_ref.seeded_params(int(z["seed"]), is_blender, is_6dof, head_scale=head_scale)
    net = _make_net(is_blender, is_6dof, baseline, params)
    x = torch.tensor(z["x"]).cuda()
    t = torch.tensor(z["t"]).cuda()
    uniform = bool(np.all(z["t"] == z["t"][0]))
    if uniform:
        t = t[:1].expand(x.shape[0], 1)  # the fid.expand(N,1) pattern -> stride 0
    with torch.no_grad():
        d_xyz, rot, sc = net(x, t)
    assert _rel(d_xyz.cpu(), z["d_xyz"]) < TOL
    if baseline:
        assert _rel(rot.cpu(), z["rot"]) < TOL
        assert _rel(sc.cpu(), z["scale"]) < TOL
    else:
        assert rot == 0 and sc == 0


@pytest.mark.parametrize("N", [1, 127, 128, 129, 20000, 100000])
def test_forward_sizes_vs_oracle(N):
    params = mlp_ref.seeded_params(3)
    net = _make_net(True, False, True, params)
    rng = np.random.default_rng(N)
    x = rng.uniform(-1.3, 1.3, size=(N, 3)).astype(np.float32)
    with torch.no_grad():
        d_xyz, rot, sc = net(torch.tensor(x).cuda(), torch.full((1, 1), 0.37).cuda().expand(N, 1))
    ref = mlp_ref.DeformRef(params, dtype=np.float32 if N > 30000 else np.float64)
    sel = slice(None) if N <= 30000 else rng.choice(N, 20000, replace=False)
    rd, rr, rs = ref.forward(x[sel], np.array([[0.37]]))
    assert _rel(d_xyz.cpu().numpy()[sel], rd) < TOL
    assert _rel(rot.cpu().numpy()[sel], rr) < TOL
    assert _rel(sc.cpu().numpy()[sel], rs) < TOL


def test_state_dict_keys_and_param_count():
    deform = pkg("deform")
    net = deform.DeformNetworkBaseline(is_blender=True)
    assert sum(p.numel() for p in net.parameters()) == 522280
    assert set(net.state_dict().keys()) == set(mlp_ref.param_shapes(True, False).keys())


@pytest.mark.parametrize("fname,is_blender,is_6dof,baseline,head_scale", CASES[:4])
def test_backward_matches_reference_golden(fname, is_blender, is_6dof, baseline, head_scale):
    z = np.load(os.path.join(GOLD, fname))
    params = mlp_ref.seeded_params(int(z["seed"]), is_blender, is_6dof, head_scale=head_scale)
    net = _make_net(is_blender, is_6dof, baseline, params)
    x = torch.tensor(z["x"]).cuda()
    t = torch.tensor(z["t"]).cuda()
    if bool(np.all(z["t"] == z["t"][0])):
        t = t[:1].expand(x.shape[0], 1)
    d_xyz, rot, sc = net(x, t)
    loss = (d_xyz * torch.tensor(z["g_dxyz"]).cuda()).sum()
    if baseline:
        loss = loss + (rot * torch.tensor(z["g_rot"]).cuda()).sum() + (sc * torch.tensor(z["g_scale"]).cuda()).sum()
    loss.backward()
    ref = mlp_ref.DeformRef(params, is_blender, is_6dof, baseline, dtype=np.float64)
    ref.forward(z["x"], z["t"])
    g = ref.backward(z["g_dxyz"], z["g_rot"] if baseline else None, z["g_scale"] if baseline else None)
    worst = {}
    for k, p in net.named_parameters():
        if k not in g:
            assert p.grad is None or torch.all(p.grad == 0), k
            continue
        assert p.grad is not None, k
        worst[k] = _rel(p.grad.cpu(), g[k])
    bad = {k: v for k, v in worst.items() if not v < TOL}
    assert not bad, bad


def test_backward_large_vs_oracle():
    N = 20000
    params = mlp_ref.seeded_params(5)
    net = _make_net(True, False, True, params)
    rng = np.random.default_rng(1)
    x = rng.uniform(-1.3, 1.3, size=(N, 3)).astype(np.float32)
    gd, gr, gs = (rng.standard_normal((N, k)).astype(np.float32) for k in (3, 4, 3))
    d_xyz, rot, sc = net(torch.tensor(x).cuda(), torch.full((1, 1), 0.5).cuda().expand(N, 1))
    ((d_xyz * torch.tensor(gd).cuda()).sum() + (rot * torch.tensor(gr).cuda()).sum()
     + (sc * torch.tensor(gs).cuda()).sum()).backward()
    ref = mlp_ref.DeformRef(params, dtype=np.float64)
    ref.forward(x, np.array([[0.5]]))
    g = ref.backward(gd, gr, gs)
    bad = {k: _rel(p.grad.cpu(), g[k]) for k, p in net.named_parameters() if not _rel(p.grad.cpu(), g[k]) < TOL}
    assert not bad, bad
