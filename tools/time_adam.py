"""Optimizer step at C2 size (6 Gaussian groups of 100k points + the 522k-parameter MLP): torch.optim.Adam (the
reference's construction) vs the fused single-launch step (csrc/adam.cu).  CUDA events, GPU box."""
import importlib, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
optim = importlib.import_module("deformable-3d-gaussians_b200.optim")
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
dev = torch.device("cuda", 0)


def make():
    shapes = [(N, 3), (N, 1, 3), (N, 15, 3), (N, 1), (N, 3), (N, 4)]
    groups = [{"params": [torch.randn(*s, device=dev).requires_grad_(True)], "lr": 1e-3, "name": str(i)}
              for i, s in enumerate(shapes)]
    net = deform.DeformNetworkBaseline(is_blender=True).to(dev)
    groups.append({"params": list(net.parameters()), "lr": 8e-4, "name": "deform"})
    for g in groups:
        for p in g["params"]:
            p.grad = torch.randn_like(p)
    return groups


def timeit(opt, n=30):
    for _ in range(5):
        opt.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        opt.step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

t_ref = timeit(torch.optim.Adam(make(), lr=0.0, eps=1e-15))
t_ours = timeit(optim.Adam(make(), lr=0.0, eps=1e-15))
n = sum(p.numel() for g in make() for p in g["params"])
print("N=%d (%d parameters)  torch.optim.Adam %.3f ms/step | fused %.3f ms/step (%.0f GB/s of 28 B/param)"
      % (N, n, t_ref, t_ours, 28 * n / t_ours / 1e6))
