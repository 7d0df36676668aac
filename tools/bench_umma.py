"""tcgen05 kind::tf32 instruction throughput: K-major vs MN-major operands, N = 256 / 96 / 16 (GPU box)."""
import ctypes as C, importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
_lib = importlib.import_module("deformable-3d-gaussians_b200._lib")
lib = _lib.load()

def desc(lbo, sbo, layout, addr=0):
    return (addr >> 4) | ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (1 << 46) | (layout << 61)

def idesc(M, N, a_mn, b_mn):
    return (1 << 4) | (2 << 7) | (2 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)

def timeit(name, ads, bds, idsc, N, reps=2000):
    a = torch.zeros(98304 // 4, device="cuda"); b = torch.zeros(98304 // 4, device="cuda")
    D = torch.zeros(128, N, device="cuda")
    adt = torch.tensor(ads, dtype=torch.int64).cuda(); bdt = torch.tensor(bds, dtype=torch.int64).cuda()
    def go(r):
        _lib.check(lib.d3g_selftest_umma_raw(_lib.ptr(a), a.numel() * 4, _lib.ptr(b), b.numel() * 4, len(ads), _lib.ptr(adt),
                                             _lib.ptr(bdt), C.c_uint32(idsc), N, _lib.ptr(D), r, C.c_void_p(0)))
    go(10)
    ts = []
    for r in (reps, 2 * reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); go(r); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    per = (ts[1] - ts[0]) * 1e-3 / (reps * len(ads))
    print("%-40s %7.1f ns/MMA  = %6.0f clk @1.965GHz   (%.2f TFLOP/s per SM)" % (name, per * 1e9, per * 1.965e9, 2 * 128 * N * 8 / per / 1e12))

k4 = range(4)
timeit("SS K-major  N=256", [desc(16, 1024, 2, 32 * k) for k in k4], [desc(16, 1024, 2, 32 * k) for k in k4], idesc(128, 256, 0, 0), 256)
timeit("SS MN-major N=256", [desc(4096, 512, 1, 1024 * k) for k in k4], [desc(4096, 512, 1, 1024 * k) for k in k4], idesc(128, 256, 1, 1), 256)
timeit("SS A MN / B K N=256", [desc(4096, 512, 1, 1024 * k) for k in k4], [desc(16, 1024, 2, 32 * k) for k in k4], idesc(128, 256, 1, 0), 256)
timeit("SS A K / B MN N=256", [desc(16, 1024, 2, 32 * k) for k in k4], [desc(4096, 512, 1, 1024 * k) for k in k4], idesc(128, 256, 0, 1), 256)
timeit("SS MN-major N=96", [desc(4096, 512, 1, 1024 * k) for k in k4], [desc(4096, 512, 1, 1024 * k) for k in k4], idesc(128, 96, 1, 1), 96)
timeit("SS A MN / ones K N=16", [desc(4096, 512, 1, 1024 * k) for k in k4], [desc(16, 1024, 2, 0) for k in k4], idesc(128, 16, 1, 0), 16)
timeit("SS K-major N=16", [desc(16, 1024, 2, 32 * k) for k in k4], [desc(16, 1024, 2, 32 * k) for k in k4], idesc(128, 16, 0, 0), 16)
timeit("SS K-major N=128", [desc(16, 1024, 2, 32 * k) for k in k4], [desc(16, 1024, 2, 32 * k) for k in k4], idesc(128, 128, 0, 0), 128)

# ---- in-kernel clocks per MMA: A from shared vs tensor memory, with/without epilogue-like TMEM traffic ----
def clk(mode, ctas, reps=4000):
    out = torch.zeros(ctas, dtype=torch.int64, device="cuda")
    _lib.check(lib.d3g_selftest_umma_clk(mode, reps, ctas, _lib.ptr(out), C.c_void_p(0)))
    torch.cuda.synchronize()
    v = out.cpu().double() / (4 * reps)
    return v.mean().item(), v.max().item()

for ctas in (148,):
    for mode, name in ((0, "SS"), (1, "TS"), (3, "SS + tcgen05.ld/st traffic"), (2, "TS + tcgen05.ld/st traffic"), (4, "TS + 8 warps polling mbarrier"), (1 + 8, "TS, commit every 4"), (1 + 16, "TS, non-zero data"), (1 + 8 + 16, "TS, commit every 4, non-zero data"), (2 + 8 + 16, "TS + ld/st traffic, commit/4, non-zero")):
        clk(mode, ctas, 100)
        m, mx = clk(mode, ctas)
        print("%-28s ctas=%3d: %.1f clk/MMA mean, %.1f max" % (name, ctas, m, mx))
