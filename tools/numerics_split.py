"""GPU experiment: how accurate is a split-operand tensor-core product?  z = x @ W^T with K = 256 (one trunk layer),
evaluated (a) single-pass TF32, (b) 3xTF32 (hi/lo split of both operands, three products), (c) fp16 hi/lo split with
fp32 accumulation (three products, weights pre-scaled by 2^8), each against an fp64 reference.  Prints the relative
error of z and how many ReLU decisions (z + b > 0) differ from fp64.  Also measures the dense TF32 peak."""
import json, sys
import torch

torch.manual_seed(0)
dev = "cuda"
N, K, M = 20000, 256, 256
x = torch.relu(torch.randn(N, K, device=dev)) * 0.7
W = (torch.rand(M, K, device=dev) * 2 - 1) / 16
b = (torch.rand(M, device=dev) * 2 - 1) / 16
ref = x.double() @ W.double().T + b.double()


def report(name, z):
    z = z.double() + b.double()
    err = (z - ref).abs()
    flips = int(((z > 0) != (ref > 0)).sum())
    out = dict(name=name, max_rel=float(err.max() / ref.abs().max()), rms_rel=float(err.pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()),
               flips=flips, units=N * M)
    print(json.dumps(out))
    return out


def tf32_round(a):
    return ((a.view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


res = []
torch.backends.cuda.matmul.allow_tf32 = False
res.append(report("fp32 cublas", x @ W.T))
torch.backends.cuda.matmul.allow_tf32 = True
res.append(report("tf32 single", x @ W.T))
xh, wh = tf32_round(x), tf32_round(W)
xl, wl = tf32_round(x - xh), tf32_round(W - wh)
res.append(report("3xtf32 (3 gemms, fp32 adds)", xh @ wh.T + (xl @ wh.T + xh @ wl.T)))
res.append(report("3xtf32 (one K=768 gemm)", torch.cat([xl, xh, xh], 1) @ torch.cat([wh, wl, wh], 1).T))
S = 256.0
xh16 = x.half(); xl16 = (x - xh16.float()).half()
Ws = W * S
wh16 = Ws.half(); wl16 = (Ws - wh16.float()).half()
mm = lambda a, c: torch.mm(a, c.T.contiguous(), out_dtype=torch.float32)
res.append(report("fp16 hi only", mm(xh16, wh16) / S))
res.append(report("fp16x3 (3 gemms, fp32 adds)", (mm(xh16, wh16) + (mm(xl16, wh16) + mm(xh16, wl16))) / S))
res.append(report("fp16x3 (one K=768 gemm)", mm(torch.cat([xl16, xh16, xh16], 1), torch.cat([wh16, wl16, wh16], 1)) / S))
res.append(report("fp16x2 (x hi only, W hi+lo)", (mm(xh16, wh16) + mm(xh16, wl16)) / S))

# dense TF32 peak, 8192^3, best of 10 and sustained (as MEASURED_PEAKS.json does for bf16)
a = torch.randn(8192, 8192, device=dev); c = torch.randn(8192, 8192, device=dev)
for _ in range(3): a @ c
best = 1e9
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); a @ c; e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
n = 0
import time
t0 = time.time()
while time.time() - t0 < 3.0:
    for _ in range(10): a @ c
    n += 10
    torch.cuda.synchronize()
e1.record(); torch.cuda.synchronize()
peak = dict(tf32_tflops=2 * 8192 ** 3 / (best * 1e-3) / 1e12, tf32_tflops_sustained=2 * 8192 ** 3 * n / (e0.elapsed_time(e1) * 1e-3) / 1e12,
            how="torch.matmul fp32 8192^3 with allow_tf32=True: best of 10 (burst), back to back for 3 s (sustained)")
print(json.dumps(peak))
json.dump(dict(split=res, peak=peak), open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/numerics_split.json", "w"), indent=1)
