import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mlp_ref
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")
ode = importlib.import_module("deformable-3d-gaussians_b200.ode")
params = {k: v for k, v in mlp_ref.seeded_params(31, True, False).items() if not k.startswith(("gaussian_rotation", "gaussian_scaling"))}
N = 96
x0 = np.random.default_rng(5).uniform(-1, 1, size=(N, 3)).astype(np.float32)
full = dict(mlp_ref.seeded_params(31, True, False)); full.update(params)
ref = mlp_ref.DeformRef(full, True, False, False, dtype=np.float64)
f = lambda t, x: ref.forward(x, np.array([[t]]))[0] * 3.0
t0 = time.time()
x = x0.astype(np.float64); t = 0.0; want = [x.copy()]
for t1 in (0.35, 0.8):
    n = 30; h = (t1 - t) / n
    for _ in range(n):
        k1 = f(t, x); k2 = f(t + h / 2, x + h / 2 * k1); k3 = f(t + h / 2, x + h / 2 * k2); k4 = f(t + h, x + h * k3)
        x = x + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4); t += h
    want.append(x.copy())
want = np.stack(want)
print("oracle rk4 %.1fs moved %.3f" % (time.time() - t0, np.abs(want[-1] - want[0]).max()))
for rtol, atol in ((1e-3, 1e-4), (1e-4, 1e-5), (1e-5, 1e-6), (1e-6, 1e-7)):
    model = deform.DeformModelODE(is_blender=True, rtol=rtol, atol=atol, output_scale=3.0)
    model.deform.load_state_dict({k: torch.tensor(v) for k, v in params.items()})
    ode.STATS.update(steps=0, rejected=0)
    t0 = time.time()
    xyz = torch.tensor(x0).cuda().requires_grad_(True)
    pos, _, _ = model.step(xyz, [0.0, 0.35, 0.8])
    torch.cuda.synchronize(); t1 = time.time()
    fs = dict(ode.STATS)
    pos.sum().backward()
    torch.cuda.synchronize()
    print("rtol %g: err %.2e  fwd steps %s %.2fs | bwd steps %d rej %d %.2fs" % (rtol, np.abs(pos.detach().cpu().numpy() - want).max(), fs, t1 - t0,
          ode.STATS["steps"] - fs["steps"], ode.STATS["rejected"] - fs["rejected"], time.time() - t1))
