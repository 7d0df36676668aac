"""Same-box CUDA baseline for the MLP half of the path: the reference's DeformNetworkBaseline (utils/time_utils.py:
56-127, is_blender=True) restated with plain torch ops (nn.Linear -> cuBLAS, ReLU, cat: what the reference launches),
forward + backward at N points on the current GPU, fp32 (allow_tf32=False, torch's default and the reference's
setting) and with TF32 cuBLAS allowed.  /root/reference is not on the GPU box, hence the restatement; its outputs are
checked against the fused kernels' in the same run.  Writes one JSON object."""
import importlib, json, os, sys
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def embed(x, L):
    out = [x]
    for f in range(L):
        out += [torch.sin(x * 2.0 ** f), torch.cos(x * 2.0 ** f)]
    return torch.cat(out, -1)


class TorchDeform(nn.Module):
    def __init__(self):
        super().__init__()
        self.timenet = nn.Sequential(nn.Linear(13, 256), nn.ReLU(inplace=True), nn.Linear(256, 30))
        self.linear = nn.ModuleList([nn.Linear(93, 256)] + [nn.Linear(256, 256) if i != 4 else nn.Linear(349, 256) for i in range(7)])
        self.gaussian_warp, self.gaussian_rotation, self.gaussian_scaling = nn.Linear(256, 3), nn.Linear(256, 4), nn.Linear(256, 3)

    def forward(self, x, t):
        te = self.timenet(embed(t, 6))
        xe = embed(x, 10)
        h = torch.cat([xe, te], -1)
        for i, l in enumerate(self.linear):
            h = torch.relu(l(h))
            if i == 4:
                h = torch.cat([xe, te, h], -1)
        return self.gaussian_warp(h), self.gaussian_rotation(h), self.gaussian_scaling(h)


def timeit(fn, warm=5, n=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
    out_path = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/torch_mlp_baseline.json"
    torch.manual_seed(0)
    deform = importlib.import_module("deformable-3d-gaussians_b200.deform")
    ours = deform.DeformNetworkBaseline(is_blender=True).cuda()
    ref = TorchDeform().cuda()
    ref.load_state_dict(ours.state_dict())
    x = torch.rand(N, 3, device="cuda") * 2.6 - 1.3
    t = torch.full((1, 1), 0.37, device="cuda").expand(N, 1)
    cot = [torch.randn(N, k, device="cuda") for k in (3, 4, 3)]

    def step(net):
        for p in net.parameters():
            p.grad = None
        o = net(x, t)
        torch.autograd.backward(o, cot)

    res = dict(N=N, gpu=torch.cuda.get_device_name(0))
    torch.backends.cuda.matmul.allow_tf32 = False
    res["torch_fp32_ms"] = timeit(lambda: step(ref))
    with torch.no_grad():
        res["torch_fp32_fwd_ms"] = timeit(lambda: ref(x, t))
    torch.backends.cuda.matmul.allow_tf32 = True
    res["torch_tf32_ms"] = timeit(lambda: step(ref))
    torch.backends.cuda.matmul.allow_tf32 = False
    res["fused_ms"] = timeit(lambda: step(ours))
    with torch.no_grad():
        res["fused_fwd_ms"] = timeit(lambda: ours(x, t))
        a, b = ours(x, t)[0], ref(x, t)[0]
        res["fused_vs_torch_fp32_dxyz_rel"] = float((a - b).abs().max() / b.abs().max())
    res["flop_per_point"] = 3.049e6
    for k in ("torch_fp32_ms", "torch_tf32_ms", "fused_ms"):
        res[k.replace("_ms", "_tflops")] = 3.049e6 * N / (res[k] * 1e-3) / 1e12
    print(json.dumps(res))
    json.dump(res, open(out_path, "w"), indent=1)


if __name__ == "__main__":
    main()
