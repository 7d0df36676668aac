"""GPU: the ODE right-hand side DeformNetworkODE.forward(t, x) (utils/time_utils.py:392-438) on the fused kernels
against goldens written by the reference's own class (tests/golden/make_golden_ode.py), and DeformModelODE.step
(scene/deform_model.py:183-207) — adaptive dopri5 + adjoint over the fused kernels — against a fixed-step fp64
Runge-Kutta integration of the numpy oracle and against back-propagation through that integration."""
import os

import numpy as np
import pytest
import torch

from conftest import pkg
from oracle import mlp_ref

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def _relF(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def _ode_params(seed, is_blender):
    return {k: v for k, v in mlp_ref.seeded_params(seed, is_blender, False).items()
            if not k.startswith(("gaussian_rotation", "gaussian_scaling"))}


@pytest.mark.parametrize("fname,is_blender", [("ode_func_blender.npz", True), ("ode_func_real.npz", False)])
def test_ode_func_matches_reference_golden(fname, is_blender):
    deform = pkg("deform")
    z = np.load(os.path.join(GOLD, fname))
    params = _ode_params(int(z["seed"]), is_blender)
    net = deform.DeformNetworkODE(is_blender=is_blender, output_scale=float(z["output_scale"]))
    assert set(net.state_dict().keys()) == set(params.keys())
    net.load_state_dict({k: torch.tensor(v) for k, v in params.items()})
    net = net.cuda()
    x = torch.tensor(z["x"]).cuda().requires_grad_(True)
    t = torch.tensor(float(z["t"])).cuda().requires_grad_(True)   # 0-dim, as the solver passes it
    f = net(t, x)
    assert f.shape == x.shape
    assert _rel(f.detach().cpu(), z["f"]) < 2e-5
    (f * torch.tensor(z["g"]).cuda()).sum().backward()
    assert _relF(x.grad.cpu(), z["grad_x"]) < 1e-3 and _rel(x.grad.cpu(), z["grad_x"]) < 1.25e-3
    # dL/dt of a shared t is the sum of N cancelling per-row terms: judge it against their natural scale
    assert abs(float(t.grad) - float(z["grad_t"])) < 1e-3 * max(abs(float(z["grad_t"])), 1.0)
    for k, p in net.named_parameters():
        want = z["grad/" + k]
        got = p.grad.cpu().numpy()[:want.shape[0]]
        assert _relF(got, want) < 1e-3 and _rel(got, want) < 1.25e-3, k
    net.check_status()


def _torch_func(params, scale):
    """Plain-torch fp64 CPU copy of DeformNetworkODE(is_blender=True) built from `params` (leaf tensors returned)."""
    P = {k: torch.tensor(v, dtype=torch.float64, requires_grad=True) for k, v in params.items()}

    def embed(v, L):
        out = [v]
        for f in range(L):
            out += [torch.sin(v * 2.0 ** f), torch.cos(v * 2.0 ** f)]
        return torch.cat(out, -1)

    class F(torch.nn.Module):
        def parameters(self):
            return list(P.values())

        def forward(self, t, x):
            te = embed(t.expand(x.shape[0], 1), 6)
            te = torch.relu(te @ P["timenet.0.weight"].T + P["timenet.0.bias"]) @ P["timenet.2.weight"].T + P["timenet.2.bias"]
            xe = embed(x, 10)
            h = torch.cat([xe, te], -1)
            for i in range(8):
                h = torch.relu(h @ P["linear.%d.weight" % i].T + P["linear.%d.bias" % i])
                if i == 4:
                    h = torch.cat([xe, te, h], -1)
            return (h @ P["gaussian_warp.weight"].T + P["gaussian_warp.bias"]) * scale

    return F(), P


def test_deform_model_ode_step_matches_the_same_solver_on_a_torch_fp64_copy():
    """DeformModelODE.step at the tolerances of arguments/ode_batch10.py (rtol 1e-4, atol 1e-5):
    positions and adjoint gradients against the SAME solver driving a plain-torch fp64 copy of the network on the
    CPU — what remains is the fused kernels' arithmetic inside the solver loop.  (The solver itself is checked on a
    system with a closed-form solution in tests/test_ode_solver.py.)"""
    deform, ode = pkg("deform"), pkg("ode")
    N, times = 64, [0.0, 0.3, 0.6]
    params = _ode_params(33, True)
    model = deform.DeformModelODE(is_blender=True, rtol=1e-4, atol=1e-5, output_scale=0.5)
    model.deform.load_state_dict({k: torch.tensor(v) for k, v in params.items()})
    rng = np.random.default_rng(6)
    x0 = rng.uniform(-1.0, 1.0, size=(N, 3)).astype(np.float32)
    cot = rng.standard_normal((3, N, 3))
    xyz = torch.tensor(x0).cuda().requires_grad_(True)
    ode.STATS.update(steps=0, rejected=0)
    pos, rot, sc = model.step(xyz, times)
    assert pos.shape == (3, N, 3) and rot.shape == (3, N, 4) and sc.shape == (3, N, 3)
    assert float(rot.abs().max()) == 0.0 and float(sc.abs().max()) == 0.0
    (pos * torch.tensor(cot, dtype=torch.float32).cuda()).sum().backward()
    assert ode.STATS["steps"] < 400, ode.STATS          # the controller is not fighting rounding noise

    func, P = _torch_func(params, 0.5)
    x = torch.tensor(x0, dtype=torch.float64, requires_grad=True)
    want = ode.odeint_adjoint(func, x, torch.tensor(times, dtype=torch.float64), rtol=1e-4, atol=1e-5)
    (want * torch.tensor(cot)).sum().backward()
    moved = float((want[-1] - want[0]).detach().abs().max())
    assert moved > 0.01                                                   # the flow actually moves the points
    assert float((pos.detach().cpu().double() - want.detach()).abs().max()) < 1e-3 * moved
    print("ode parity: pos %.2e dx %.2e steps %s" % (float((pos.detach().cpu().double() - want.detach()).abs().max()) / moved,
                                                      _relF(xyz.grad.cpu(), x.grad), ode.STATS))
    assert _relF(xyz.grad.cpu(), x.grad) < 1e-2
    errs = {k: _relF(p.grad.cpu(), P[k].grad) for k, p in model.deform.named_parameters()}
    # (two adaptive solves whose right-hand sides differ in the last bits accept / reject slightly different steps; a
    # single vjp of the same network meets the 1e-3 bar in test_ode_func_matches_reference_golden)
    bad = {k: e for k, e in errs.items() if not e < 5e-2}
    assert not bad, bad
    # the reference's validation shortcut: a single time 0 returns the canonical positions untouched
    p0, r0, s0 = model.step(xyz, [0.0])
    assert p0 is xyz and r0.shape == (N, 4) and s0.shape == (N, 3)
    model.deform.check_status()
