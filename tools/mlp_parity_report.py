"""Parity report of the fused MLP against the fp64 oracle WITHOUT handing the oracle the GPU's ReLU pattern:
per fixture and at N = 20k: ReLU decisions that differ, forward error, and for every parameter gradient
max|a-b|/max|b| and ||a-b||/||b||.  Output: one JSON object per case."""
import importlib, json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mlp_ref
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")
GOLD = os.path.join(ROOT, "tests", "golden")
rel = lambda a, b: float(np.abs(np.asarray(a, np.float64) - b).max() / max(np.abs(b).max(), 1e-30))
relF = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / max(np.linalg.norm(b), 1e-30))


def run(name, params, is_blender, is_6dof, baseline, x, t, gd, gr, gs):
    cls = deform.DeformNetworkBaseline if baseline else deform.DeformNetwork
    net = cls(is_blender=is_blender, is_6dof=is_6dof)
    net.load_state_dict({k: torch.tensor(v) for k, v in params.items()})
    net = net.cuda(); net.debug = True
    xt = torch.tensor(x).cuda().requires_grad_(True)
    tt = torch.tensor(t).cuda()
    if np.all(t == t[0]):
        tt = tt[:1].expand(x.shape[0], 1)
    outs = net(xt, tt)
    loss = (outs[0] * torch.tensor(gd).cuda()).sum()
    if baseline:
        loss = loss + (outs[1] * torch.tensor(gr).cuda()).sum() + (outs[2] * torch.tensor(gs).cuda()).sum()
    loss.backward()
    masks = net.read_activation_masks(deform._FusedDeform.last_ws, x.shape[0])
    ref = mlp_ref.DeformRef(params, is_blender, is_6dof, baseline, dtype=np.float64)
    ro = ref.forward(x, t)
    g = ref.backward(gd, gr if baseline else None, gs if baseline else None, want_input_grads=True)
    flips = sum(int((masks[l] != (ref.cache["z"][l] > 0)).sum()) for l in range(8))
    out = dict(case=name, rows=int(x.shape[0]), relu_flips=flips, fwd_rel=rel(outs[0].detach().cpu().numpy(), ro[0]))
    worst, worstF = ("", 0.0), ("", 0.0)
    for k, p in net.named_parameters():
        if k not in g or p.grad is None:
            continue
        e, f = rel(p.grad.cpu().numpy(), g[k]), relF(p.grad.cpu().numpy(), g[k])
        if e > worst[1]: worst = (k, e)
        if f > worstF[1]: worstF = (k, f)
    out["worst_max_rel"], out["worst_l2_rel"] = worst, worstF
    out["dx_max_rel"] = rel(xt.grad.cpu().numpy(), g["_x"])
    print(json.dumps(out))
    return out


res = []
CASES = [("mlp_blender.npz", True, False, True, 1.0), ("mlp_blender_dn.npz", True, False, False, 1.0),
         ("mlp_real.npz", False, False, True, 1.0), ("mlp_6dof.npz", True, True, True, 0.05),
         ("mlp_blender_rowt.npz", True, False, True, 1.0)]
for fname, b, s, base, hs in CASES:
    z = np.load(os.path.join(GOLD, fname))
    params = mlp_ref.seeded_params(int(z["seed"]), b, s, head_scale=hs)
    res.append(run(fname, params, b, s, base, z["x"], z["t"], z["g_dxyz"], z["g_rot"] if base else None, z["g_scale"] if base else None))
for N, seed in ((20000, 5), (20000, 6), (50000, 7)):
    params = mlp_ref.seeded_params(seed)
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1.3, 1.3, size=(N, 3)).astype(np.float32)
    gd, gr, gs = (rng.standard_normal((N, k)).astype(np.float32) for k in (3, 4, 3))
    res.append(run("N=%d seed %d" % (N, seed), params, True, False, True, x, np.full((N, 1), 0.5, np.float32), gd, gr, gs))
json.dump(res, open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/mlp_parity.json", "w"), indent=1)
