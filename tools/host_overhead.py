"""Where does the host time of one fused-MLP call go?  cProfile over 300 no_grad forwards + wall clock per call."""
import cProfile, importlib, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")
N = 1000
net = deform.DeformNetworkBaseline(is_blender=True).cuda()
x = torch.rand(N, 3, device="cuda")
t = torch.full((1, 1), 0.37, device="cuda").expand(N, 1)
with torch.no_grad():
    for _ in range(20):
        net(x, t)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(300):
        net(x, t)
    torch.cuda.synchronize()
    print("wall per call: %.1f us" % ((time.perf_counter() - t0) / 300 * 1e6))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(300):
        net(x, t)
    pr.disable()
    torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
