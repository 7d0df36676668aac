"""Per-parameter gradient error report of the fused MLP backward vs the numpy oracle (GPU box)."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import mlp_ref
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")

def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)

def run(N, is_blender, is_6dof, baseline, head_scale=1.0, seed=5):
    params = mlp_ref.seeded_params(seed, is_blender, is_6dof, head_scale=head_scale)
    cls = deform.DeformNetworkBaseline if baseline else deform.DeformNetwork
    net = cls(is_blender=is_blender, is_6dof=is_6dof)
    net.load_state_dict({k: torch.tensor(v) for k, v in params.items()})
    net.debug = True
    net = net.cuda()
    rng = np.random.default_rng(1)
    x = rng.uniform(-1.3, 1.3, size=(N, 3)).astype(np.float32)
    gd = rng.standard_normal((N, 4, 4) if is_6dof else (N, 3)).astype(np.float32)
    gr, gs = (rng.standard_normal((N, k)).astype(np.float32) for k in (4, 3))
    out = net(torch.tensor(x).cuda(), torch.full((1, 1), 0.5).cuda().expand(N, 1))
    loss = (out[0] * torch.tensor(gd).cuda()).sum()
    if baseline:
        loss = loss + (out[1] * torch.tensor(gr).cuda()).sum() + (out[2] * torch.tensor(gs).cuda()).sum()
    loss.backward()
    ref = mlp_ref.DeformRef(params, is_blender, is_6dof, baseline, dtype=np.float64)
    ref.forward(x, np.array([[0.5]]))
    g = ref.backward(gd, gr if baseline else None, gs if baseline else None)
    print("== N=%d blender=%d 6dof=%d baseline=%d" % (N, is_blender, is_6dof, baseline))
    for k, p in net.named_parameters():
        if k in g and p.grad is not None:
            print("  %-28s rel %.3e  |ref|max %.3e |got|max %.3e" % (k, rel(p.grad.cpu(), g[k]), np.abs(g[k]).max(), p.grad.abs().max().item()))
        else:
            print("  %-28s ref %s got %s" % (k, k in g, p.grad is not None))

if __name__ == "__main__":
    run(128, True, False, True)
    run(1000, True, False, True)
    run(20000, True, False, True)
    run(1000, False, False, True)
    run(1000, True, True, True, head_scale=0.05)
