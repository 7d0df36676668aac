"""Time the fused MLP kernels alone (CUDA events): inference forward (nothing saved), training forward, backward."""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
deform = importlib.import_module("deformable-3d-gaussians_b200.deform")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
torch.manual_seed(0)
net = deform.DeformNetworkBaseline(is_blender=True).cuda()
x = (torch.rand(N, 3, device="cuda") * 2.6 - 1.3)
t = torch.full((1, 1), 0.37, device="cuda").expand(N, 1)


def timeit(fn, n=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def infer():
    with torch.no_grad():
        net(x, t)


outs = None
def train_fwd():
    global outs
    outs = net(x, t)

cot = [torch.randn(N, 3, device="cuda"), torch.randn(N, 4, device="cuda"), torch.randn(N, 3, device="cuda")]
def train_fwd_bwd():
    o = net(x, t)
    torch.autograd.backward(o, cot)

print("N=%d  infer fwd %.3f ms | train fwd %.3f ms | train fwd+bwd %.3f ms" % (N, timeit(infer), timeit(train_fwd), timeit(train_fwd_bwd)))
